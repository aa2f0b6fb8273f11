// ref_harness.cpp — TEST INFRASTRUCTURE ONLY.
//
// A thin extern "C" shim over the UNMODIFIED reference classes (compiled from the sources where they lie under
// /root/reference by oracle/Makefile, output only into oracle/_ref/).  It lets tests/ and bench.py's cpu_baseline /
// `--impl reference` leg call the reference's own Packing / Interaction / NeighbourGrid / Simulation through ctypes.
// Nothing in the product path (rampack_b200/) may link or load this.
//
// The shim contains no algorithm of its own: every function forwards to a reference method (cited inline).

#include <memory>
#include <vector>
#include <sstream>
#include <cstring>
#include <cstdint>
#include <limits>

#include "core/Packing.h"
#include "core/observables/NematicOrder.h"
#include "core/observables/SmecticOrder.h"
#include "core/observables/correlation/PairDensityCorrelation.h"
#include "core/observables/correlation/RadialEnumerator.h"
#include <cstdio>
#include "core/PeriodicBoundaryConditions.h"
#include "core/NeighbourGrid.h"
#include "core/Simulation.h"
#include "core/ObservablesCollector.h"
#include "core/observables/NumberDensity.h"
#include "core/shapes/SphereTraits.h"
#include "core/shapes/SpherocylinderTraits.h"
#include "core/shapes/PolysphereTraits.h"
#include "core/shapes/PolyspherocylinderTraits.h"
#include "core/shapes/CompoundShapeTraits.h"
#include "core/shapes/XenoCollideTraits.h"
#include "core/shapes/SmoothWedgeTraits.h"
#include "core/shapes/PolyhedralWedgeTraits.h"
#include "geometry/xenocollide/XCPrimitives.h"
#include "geometry/xenocollide/XCOperations.h"
#include "core/interactions/LennardJonesInteraction.h"
#include "core/interactions/RepulsiveLennardJonesInteraction.h"
#include "core/interactions/SquareInverseCoreInteraction.h"
#include "core/move_samplers/RototranslationSampler.h"
#include "core/volume_scalers/TriclinicDeltaScaler.h"
#include "core/volume_scalers/TriclinicAdapter.h"
#include "core/volume_scalers/DeltaVolumeScaler.h"
#include "core/volume_scalers/AnisotropicVolumeScaler.h"
#include "core/volume_scalers/LinearScalingFactorSampler.h"
#include "core/volume_scalers/LogScalingFactorSampler.h"
#include "core/lattice/UnitCell.h"
#include "core/lattice/Lattice.h"
#include "core/lattice/SerialPopulator.h"
#include "utils/Logger.h"
#include "utils/OMPMacros.h"

#include "../include/rampack_b200.h"

namespace {
    // Same role as the anonymous HardcodedTranslation in src/core/Packing.cpp:23-43 (not exported by the reference)
    class FixedTranslation : public BoundaryConditions {
        Vector<3> t;
    public:
        explicit FixedTranslation(const Vector<3> &t) : t{t} { }
        void setBox(const TriclinicBox &) override { }
        [[nodiscard]] Vector<3> getCorrection(const Vector<3> &) const override { return {}; }
        [[nodiscard]] Vector<3> getTranslation(const Vector<3> &, const Vector<3> &) const override { return t; }
        [[nodiscard]] double getDistance2(const Vector<3> &p1, const Vector<3> &p2) const override {
            return (p2 + t - p1).norm2();
        }
    };

    Matrix<3, 3> mat(const double *a) { return Matrix<3, 3>(a); }
    Vector<3> vec(const double *a) { return Vector<3>{a[0], a[1], a[2]}; }

    struct Traits {
        std::shared_ptr<ShapeTraits> traits;
    };

    std::shared_ptr<CentralInteraction> makeSoft(const rp_interaction_desc *d) {
        switch (d->soft_kind) {
            case RP_SOFT_LJ:  return std::make_shared<LennardJonesInteraction>(d->epsilon, d->sigma);
            case RP_SOFT_WCA: return std::make_shared<RepulsiveLennardJonesInteraction>(d->epsilon, d->sigma);
            case RP_SOFT_SIC: return std::make_shared<SquareInverseCoreInteraction>(d->epsilon, d->sigma);
            default: return nullptr;
        }
    }

    std::vector<Shape> toShapes(const double *shapes, std::uint64_t n) {
        std::vector<Shape> out;
        out.reserve(n);
        for (std::uint64_t i = 0; i < n; i++)
            out.emplace_back(vec(shapes + 12*i), mat(shapes + 12*i + 3));
        return out;
    }

    void fromShape(const Shape &s, double *out) {
        for (int k = 0; k < 3; k++) out[k] = s.getPosition()[k];
        for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) out[3 + 3*r + c] = s.getOrientation()(r, c);
    }

    // RP_HARD_XENOCOLLIDE: the reference's OWN geometry classes instantiated from the flat program the C ABI takes
    // (include/rampack_b200.h, rp_xc_program).  Only public constructors / add() are used, in the order a
    // describe() walk of the reference tree emits them, so the objects are what XCBodyBuilder would have built.
    std::shared_ptr<AbstractXCGeometry> buildXcNode(const rp_xc_program &P, int idx,
                                                    std::vector<std::shared_ptr<AbstractXCGeometry>> &built)
    {
        const rp_xc_node &N = P.nodes[idx];
        const double *a = N.params;
        switch (N.kind) {
            case RP_XC_POINT:     return std::make_shared<XCPoint>(Vector<3>{a[0], a[1], a[2]});
            case RP_XC_SEGMENT:   return std::make_shared<XCSegment>(a[0]);
            case RP_XC_RECTANGLE: return std::make_shared<XCRectangle>(a[0], a[1]);
            case RP_XC_CUBOID:    return std::make_shared<XCCuboid>(Vector<3>{a[0], a[1], a[2]});
            case RP_XC_DISK:      return std::make_shared<XCDisk>(a[0]);
            case RP_XC_SPHERE:    return std::make_shared<XCSphere>(a[0]);
            case RP_XC_ELLIPSE:   return std::make_shared<XCEllipse>(a[0], a[1]);
            case RP_XC_ELLIPSOID: return std::make_shared<XCEllipsoid>(Vector<3>{a[0], a[1], a[2]});
            case RP_XC_FOOTBALL:  return std::make_shared<XCFootball>(a[0], a[1]);
            case RP_XC_SAUCER:    return std::make_shared<XCSaucer>(a[0], a[1]);
            case RP_XC_SMOOTH_WEDGE:
                return std::make_shared<PolymorphicXCAdapter<SmoothWedgeTraits::CollideGeometry>>(
                    SmoothWedgeTraits::CollideGeometry(a[0], a[1], a[2]));
            case RP_XC_POLYHEDRAL_WEDGE:
                return std::make_shared<PolymorphicXCAdapter<PolyhedralWedgeTraits::CollideGeometry>>(
                    PolyhedralWedgeTraits::CollideGeometry(a[0], a[1], a[2], a[3], a[4]));
            case RP_XC_SUM: {
                auto g = std::make_shared<XCSum>();
                for (int e = 0; e < N.n_entries; e++) {
                    const rp_xc_entry &E = P.entries[N.first_entry + e];
                    g->add(built.at(E.node), vec(E.pos), mat(E.rot));
                }
                return g;
            }
            case RP_XC_MAX: {
                auto g = std::make_shared<XCMax>();
                for (int e = 0; e < N.n_entries; e++) {
                    const rp_xc_entry &E = P.entries[N.first_entry + e];
                    g->add(built.at(E.node), vec(E.pos), mat(E.rot));
                }
                return g;
            }
            case RP_XC_DIFF: {
                const rp_xc_entry &E1 = P.entries[N.first_entry], &E2 = P.entries[N.first_entry + 1];
                return std::make_shared<XCDiff>(built.at(E1.node), mat(E1.rot), vec(E1.pos),
                                                built.at(E2.node), mat(E2.rot), vec(E2.pos));
            }
            default:
                throw std::runtime_error("unknown rp_xc_kind");
        }
    }

    // Same pattern as GenericXenoCollideTraits (one geometry) and SmoothWedgeTraits / PolyhedralWedgeTraits (one
    // geometry per interaction centre): the CRTP base supplies overlapBetween / getRangeRadius unchanged.
    class ProgramXCTraits : public XenoCollideTraits<ProgramXCTraits> {
        std::vector<std::shared_ptr<AbstractXCGeometry>> geometries;
        std::vector<Vector<3>> centres;
    public:
        ProgramXCTraits(std::vector<std::shared_ptr<AbstractXCGeometry>> geometries, std::vector<Vector<3>> centres)
                : XenoCollideTraits({0, 0, 1}, {1, 0, 0}, {0, 0, 0}, 1.0),
                  geometries{std::move(geometries)}, centres{std::move(centres)}
        { }
        [[nodiscard]] const AbstractXCGeometry &getCollideGeometry(std::size_t i = 0) const {
            return *this->geometries.at(i);
        }
        [[nodiscard]] std::vector<Vector<3>> getInteractionCentres() const override { return this->centres; }
    };

    std::shared_ptr<ShapeTraits> makeXcTraits(const rp_interaction_desc *d) {
        if (!d->xc) throw std::runtime_error("RP_HARD_XENOCOLLIDE without a program");
        const rp_xc_program &P = *d->xc;
        std::vector<std::shared_ptr<AbstractXCGeometry>> built;
        for (int i = 0; i < P.n_nodes; i++) built.push_back(buildXcNode(P, i, built));
        std::vector<std::shared_ptr<AbstractXCGeometry>> geoms;
        std::vector<Vector<3>> centres;
        for (int c = 0; c < (d->n_centres > 0 ? d->n_centres : 1); c++) geoms.push_back(built.at(P.root[c]));
        for (int c = 0; c < d->n_centres; c++) centres.push_back(vec(d->centres[c]));
        return std::make_shared<ProgramXCTraits>(std::move(geoms), std::move(centres));
    }
}

extern "C" {

void *rpr_traits_create(const rp_interaction_desc *d) {
    try {
        auto t = new Traits;
        auto soft = makeSoft(d);
        switch (d->hard_kind) {
            case RP_HARD_SPHERE:
                t->traits = std::make_shared<SphereTraits>(d->radius);
                break;
            case RP_HARD_SPHEROCYLINDER:
                t->traits = std::make_shared<SpherocylinderTraits>(d->length, d->radius);
                break;
            case RP_HARD_POLYSPHERE: {
                std::vector<PolysphereTraits::SphereData> data;
                for (int i = 0; i < d->n_centres; i++)
                    data.emplace_back(vec(d->centres[i]), d->centre_radius[i]);
                PolysphereTraits::PolysphereGeometry geom(std::move(data), Vector<3>{0, 0, 1}, Vector<3>{1, 0, 0});
                t->traits = std::make_shared<PolysphereTraits>(std::move(geom));
                break;
            }
            case RP_HARD_POLYSPHEROCYLINDER: {
                std::vector<PolyspherocylinderTraits::SpherocylinderData> data;
                for (int i = 0; i < d->n_centres; i++)
                    data.emplace_back(vec(d->centres[i]), vec(d->centre_half_axis[i]), d->centre_radius[i]);
                PolyspherocylinderTraits::PolyspherocylinderGeometry geom(std::move(data), Vector<3>{0, 0, 1},
                                                                          Vector<3>{1, 0, 0});
                t->traits = std::make_shared<PolyspherocylinderTraits>(std::move(geom));
                break;
            }
            case RP_HARD_XENOCOLLIDE:
                t->traits = makeXcTraits(d);
                break;
            case RP_HARD_NONE: {
                if (!soft) { delete t; return nullptr; }
                if (d->n_centres == 0) {
                    t->traits = std::make_shared<SphereTraits>(d->radius, soft);
                } else {
                    std::vector<PolysphereTraits::SphereData> data;
                    for (int i = 0; i < d->n_centres; i++)
                        data.emplace_back(vec(d->centres[i]), d->centre_radius[i]);
                    PolysphereTraits::PolysphereGeometry geom(std::move(data), Vector<3>{0, 0, 1}, Vector<3>{1, 0, 0});
                    t->traits = std::make_shared<PolysphereTraits>(std::move(geom), soft);
                }
                break;
            }
            default:
                delete t;
                return nullptr;
        }
        // hard + soft: the reference expresses it as CompoundShapeTraits(main = hard, aux = soft-only shape on the
        // same centres) (src/core/shapes/CompoundShapeTraits.h:30-34, src/frontend/modes/CasinoMode.cpp:288-289)
        if (d->hard_kind != RP_HARD_NONE && soft) {
            std::shared_ptr<ShapeTraits> aux;
            if (d->n_centres == 0) {
                aux = std::make_shared<SphereTraits>(d->radius, soft);
            } else {
                std::vector<PolysphereTraits::SphereData> data;
                for (int i = 0; i < d->n_centres; i++)
                    data.emplace_back(vec(d->centres[i]), d->centre_radius[i]);
                PolysphereTraits::PolysphereGeometry geom(std::move(data), Vector<3>{0, 0, 1}, Vector<3>{1, 0, 0});
                aux = std::make_shared<PolysphereTraits>(std::move(geom), soft);
            }
            t->traits = std::make_shared<CompoundShapeTraits>(t->traits, aux);
        }
        return t;
    } catch (const std::exception &) {
        return nullptr;
    }
}

void rpr_traits_destroy(void *t) { delete static_cast<Traits *>(t); }

double rpr_range_radius(void *t) { return static_cast<Traits *>(t)->traits->getInteraction().getRangeRadius(); }
double rpr_total_range_radius(void *t) { return static_cast<Traits *>(t)->traits->getInteraction().getTotalRangeRadius(); }

// Interaction::overlapBetween / calculateEnergyBetween on explicit pairs (src/core/Interaction.h:60-85)
int rpr_pair_test_batch(void *tp, const rp_pair *pairs, std::uint64_t n, std::uint8_t *verdict, double *energy) {
    const Interaction &inter = static_cast<Traits *>(tp)->traits->getInteraction();
    bool hard = inter.hasHardPart(), soft = inter.hasSoftPart();
    for (std::uint64_t i = 0; i < n; i++) {
        const rp_pair &p = pairs[i];
        FixedTranslation bc(vec(p.translation));
        if (verdict)
            verdict[i] = hard && inter.overlapBetween(vec(p.pos1), mat(p.orientation1), p.idx1,
                                                      vec(p.pos2), mat(p.orientation2), p.idx2, bc);
        if (energy)
            energy[i] = soft ? inter.calculateEnergyBetween(vec(p.pos1), mat(p.orientation1), p.idx1,
                                                            vec(p.pos2), mat(p.orientation2), p.idx2, bc) : 0;
    }
    return 0;
}

struct RefPacking {
    std::unique_ptr<Packing> packing;
    std::string error;
};

// Packing::Packing (src/core/Packing.cpp:46-54)
void *rpr_packing_create(void *tp, const double box[9], const double *shapes, std::uint64_t n, int moveThreads,
                         int scalingThreads)
{
    auto p = new RefPacking;
    try {
        const Interaction &inter = static_cast<Traits *>(tp)->traits->getInteraction();
        p->packing = std::make_unique<Packing>(TriclinicBox(mat(box)), toShapes(shapes, n),
                                               std::make_unique<PeriodicBoundaryConditions>(), inter,
                                               moveThreads, scalingThreads);
    } catch (const std::exception &e) {
        p->error = e.what();
        p->packing.reset();
    }
    return p;
}

int rpr_packing_ok(void *pp) { return static_cast<RefPacking *>(pp)->packing != nullptr; }
const char *rpr_packing_error(void *pp) { return static_cast<RefPacking *>(pp)->error.c_str(); }
void rpr_packing_destroy(void *pp) { delete static_cast<RefPacking *>(pp); }

void rpr_packing_get(void *pp, double box[9], double *shapes) {
    const Packing &pk = *static_cast<RefPacking *>(pp)->packing;
    const auto &dim = pk.getBox().getDimensions();
    for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) box[3*r + c] = dim(r, c);
    for (std::size_t i = 0; i < pk.size(); i++)
        fromShape(pk[i], shapes + 12*i);
}

int rpr_grid_divisions(void *pp, std::uint64_t out[3]) {
    const Packing &pk = *static_cast<RefPacking *>(pp)->packing;
    auto d = pk.getNeighbourGridCellDivisions();
    out[0] = d[0]; out[1] = d[1]; out[2] = d[2];
    return 0;
}

double rpr_average_neighbours(void *pp) {
    return static_cast<RefPacking *>(pp)->packing->getAverageNumberOfNeighbours();
}

// Packing::countTotalOverlaps (src/core/Packing.cpp:519-559)
std::uint64_t rpr_count_total_overlaps(void *pp, void *tp, int earlyExit) {
    const Interaction &inter = static_cast<Traits *>(tp)->traits->getInteraction();
    return static_cast<RefPacking *>(pp)->packing->countTotalOverlaps(inter, earlyExit != 0);
}

// Packing::getTotalEnergy (src/core/Packing.cpp:690-709)
double rpr_total_energy(void *pp, void *tp) {
    const Interaction &inter = static_cast<Traits *>(tp)->traits->getInteraction();
    return static_cast<RefPacking *>(pp)->packing->getTotalEnergy(inter);
}

// Packing::tryTranslation / tryRotation / tryMove (src/core/Packing.cpp:84-164)
int rpr_try_move(void *pp, void *tp, const rp_move *m, double *dE) {
    auto *rp = static_cast<RefPacking *>(pp);
    const Interaction &inter = static_cast<Traits *>(tp)->traits->getInteraction();
    try {
        switch (m->kind) {
            case RP_MOVE_TRANSLATION:
                *dE = rp->packing->tryTranslation(m->particle, vec(m->translation), inter);
                break;
            case RP_MOVE_ROTATION:
                *dE = rp->packing->tryRotation(m->particle, mat(m->rotation), inter);
                break;
            default:
                *dE = rp->packing->tryMove(m->particle, vec(m->translation), mat(m->rotation), inter);
        }
    } catch (const std::exception &e) {
        rp->error = e.what();
        return 1;
    }
    return 0;
}

// Packing::acceptMove (src/core/Packing.cpp:278-302) — what Simulation::tryMove always calls (Simulation.cpp:484)
void rpr_accept_move(void *pp) { static_cast<RefPacking *>(pp)->packing->acceptMove(); }

// Packing::tryScaling / revertScaling (src/core/Packing.cpp:172-208, 363-371)
int rpr_try_scaling(void *pp, void *tp, const double newBox[9], double *dE) {
    auto *rp = static_cast<RefPacking *>(pp);
    const Interaction &inter = static_cast<Traits *>(tp)->traits->getInteraction();
    try {
        *dE = rp->packing->tryScaling(TriclinicBox(mat(newBox)), inter);
    } catch (const std::exception &e) {
        rp->error = e.what();
        return 1;
    }
    return 0;
}

void rpr_revert_scaling(void *pp) { static_cast<RefPacking *>(pp)->packing->revertScaling(); }

void rpr_toggle_overlap_counting(void *pp, void *tp, int on) {
    const Interaction &inter = static_cast<Traits *>(tp)->traits->getInteraction();
    static_cast<RefPacking *>(pp)->packing->toggleOverlapCounting(on != 0, inter);
}

// Packing::toggleWalls (src/core/Packing.cpp:1081-1084)
void rpr_toggle_walls(void *pp, const int walls[3]) {
    static_cast<RefPacking *>(pp)->packing->toggleWalls({walls[0] != 0, walls[1] != 0, walls[2] != 0});
}

// NematicOrder / SmecticOrder of the reference (src/core/observables/NematicOrder.cpp, SmecticOrder.cpp) on the
// packing: out = {P2, Q11, Q12, Q13, Q22, Q23, Q33, |tau|, kx, ky, kz}, hklOut = the (h, k, l) of the maximum
int rpr_observables(void *pp, void *tp, const int hklRange[3], double out[11], int hklOut[3]) {
    try {
        const Packing &packing = *static_cast<RefPacking *>(pp)->packing;
        const ShapeTraits &traits = *static_cast<Traits *>(tp)->traits;
        NematicOrder nematic(true);
        nematic.calculate(packing, 1, 1, traits);
        auto nv = nematic.getIntervalValues();
        for (std::size_t i = 0; i < 7; i++) out[i] = nv[i];
        SmecticOrder smectic({std::size_t(hklRange[0]), std::size_t(hklRange[1]), std::size_t(hklRange[2])}, true);
        smectic.calculate(packing, 1, 1, traits);
        auto sv = smectic.getIntervalValues();
        for (std::size_t i = 0; i < 4; i++) out[7 + i] = sv[i];
        auto nominal = smectic.getNominalValues().front();
        std::sscanf(nominal.c_str(), "%d.%d.%d", &hklOut[0], &hklOut[1], &hklOut[2]);
        return 0;
    } catch (const std::exception &) {
        return 1;
    }
}

// One snapshot of the reference's PairDensityCorrelation with RadialEnumerator("o")
// (src/core/observables/correlation/PairDensityCorrelation.cpp:8-41): g(r) per bin and the pair counts, read back from
// the class's own print() (distance, value, count)
int rpr_pair_density(void *pp, void *tp, double maxR, std::uint64_t numBins, double *gOut, double *countOut) {
    try {
        const Packing &packing = *static_cast<RefPacking *>(pp)->packing;
        const ShapeTraits &traits = *static_cast<Traits *>(tp)->traits;
        PairDensityCorrelation rho(std::make_shared<RadialEnumerator>(), maxR, numBins, true);
        rho.addSnapshot(packing, 1, 1, traits);
        std::ostringstream out;
        out.precision(17);
        rho.print(out);
        std::istringstream in(out.str());
        for (std::uint64_t b = 0; b < numBins; b++) {
            double x, y, c;
            if (!(in >> x >> y >> c)) return 2;
            gOut[b] = y;
            countOut[b] = c;
        }
        return 0;
    } catch (const std::exception &) {
        return 1;
    }
}

std::uint64_t rpr_cached_overlaps(void *pp) { return static_cast<RefPacking *>(pp)->packing->getCachedNumberOfOverlaps(); }

// NeighbourGrid::positionToCellNo (src/core/NeighbourGrid.cpp:13-34) on a grid built like Packing::rebuildNeighbourGrid
int rpr_cell_numbers(const double box[9], double cellSize, const double *positions, std::uint64_t n,
                     std::uint64_t *out, std::uint64_t divisionsOut[3])
{
    try {
        NeighbourGrid ng(TriclinicBox(mat(box)), cellSize, n);
        auto d = ng.getCellDivisions();
        divisionsOut[0] = d[0]; divisionsOut[1] = d[1]; divisionsOut[2] = d[2];
        for (std::uint64_t i = 0; i < n; i++)
            out[i] = ng.positionToCellNo(vec(positions + 3*i));
    } catch (const std::exception &) {
        return 1;
    }
    return 0;
}

// PeriodicBoundaryConditions::getCorrection (src/core/PeriodicBoundaryConditions.cpp:11-22)
void rpr_pbc_correction(const double box[9], const double pos[3], double out[3]) {
    PeriodicBoundaryConditions bc;
    bc.setBox(TriclinicBox(mat(box)));
    auto c = bc.getCorrection(vec(pos));
    out[0] = c[0]; out[1] = c[1]; out[2] = c[2];
}

// TriclinicBox heights / volume / inverse (src/core/TriclinicBox.cpp:27-51)
void rpr_box_info(const double box[9], double heights[3], double *volume, double inverse[9]) {
    TriclinicBox b(mat(box));
    auto h = b.getHeights();
    heights[0] = h[0]; heights[1] = h[1]; heights[2] = h[2];
    *volume = b.getVolume();
    auto inv = b.getDimensions().inverse();
    for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) inverse[3*r + c] = inv(r, c);
}

// Matrix<3,3>::rotation (src/geometry/Matrix.tpp:437-449)
void rpr_rotation_matrix(const double axis[3], double angle, double out[9]) {
    auto R = Matrix<3, 3>::rotation(vec(axis), angle);
    for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) out[3*r + c] = R(r, c);
}

struct rpr_sim_result {
    double move_seconds;
    double scaling_seconds;
    double domain_decomposition_seconds;
    double total_seconds;
    std::uint64_t moves_attempted;
    std::uint64_t moves_accepted;
    std::uint64_t scalings_attempted;
    std::uint64_t scalings_accepted;
    double density_mean;
    double density_error;
    double ng_rebuild_seconds;
    std::uint64_t cycles;
    int threads;
    int domains[3];
};

// Simulation::integrate (src/core/Simulation.cpp:98-199) with a rototranslation sampler and the delta-triclinic /
// delta-V box scaler, exactly how test/validation_tests/SimulationTest.cpp drives it.
// box_scaler: 0 = disabled (NVT), 1 = delta_v(step) via TriclinicAdapter, 2 = delta_triclinic(step, independent)
int rpr_run_simulation(void *tp, const double box[9], const double *shapes, std::uint64_t n, std::uint64_t seed,
                       double temperature, double pressure, double transStep, double rotStep, int boxScaler,
                       double boxStep, const int domainDivisions[3], int threads, std::uint64_t thermCycles,
                       std::uint64_t avgCycles, std::uint64_t avgEvery, double boxOut[9], double *shapesOut,
                       rpr_sim_result *res, char *err, std::size_t errLen)
{
    try {
        auto traits = static_cast<Traits *>(tp)->traits;
        const Interaction &inter = traits->getInteraction();
        std::size_t numDomains = std::size_t(domainDivisions[0]) * domainDivisions[1] * domainDivisions[2];
        if (threads <= 0) threads = static_cast<int>(numDomains);
        auto packing = std::make_unique<Packing>(TriclinicBox(mat(box)), toShapes(shapes, n),
                                                 std::make_unique<PeriodicBoundaryConditions>(), inter, threads,
                                                 threads);
        std::unique_ptr<TriclinicBoxScaler> scaler;
        if (boxScaler == 1)
            scaler = std::make_unique<TriclinicAdapter>(std::make_unique<DeltaVolumeScaler>(), boxStep);
        else if (boxScaler == 2)
            scaler = std::make_unique<TriclinicDeltaScaler>(boxStep, true);

        std::array<std::size_t, 3> dd = {std::size_t(domainDivisions[0]), std::size_t(domainDivisions[1]),
                                         std::size_t(domainDivisions[2])};
        std::unique_ptr<Simulation> sim;
        if (scaler) {
            sim = std::make_unique<Simulation>(std::move(packing), transStep, rotStep, seed, std::move(scaler), dd,
                                               false);
        } else {
            Simulation::Environment env;
            std::vector<std::shared_ptr<MoveSampler>> samplers;
            samplers.push_back(std::make_shared<RototranslationSampler>(transStep, rotStep));
            env.setMoveSamplers(std::move(samplers));
            env.disableBoxScaling();
            sim = std::make_unique<Simulation>(std::move(packing), seed, std::move(env), dd, false);
        }

        auto collector = std::make_unique<ObservablesCollector>();
        collector->addObservable(std::make_unique<NumberDensity>(), ObservablesCollector::AVERAGING);
        std::ostringstream loggerStream;
        Logger logger(loggerStream);

        std::size_t total = thermCycles + avgCycles;
        sim->integrate(temperature, pressure, thermCycles, avgCycles, avgEvery == 0 ? 1 : avgEvery, total,
                       *traits, std::move(collector), {}, logger);

        res->move_seconds = sim->getMoveMicroseconds() * 1e-6;
        res->scaling_seconds = sim->getScalingMicroseconds() * 1e-6;
        res->domain_decomposition_seconds = sim->getDomainDecompositionMicroseconds() * 1e-6;
        res->total_seconds = sim->getTotalMicroseconds() * 1e-6;
        res->moves_attempted = 0;
        res->moves_accepted = 0;
        res->scalings_attempted = 0;
        res->scalings_accepted = 0;
        for (const auto &st : sim->getMovesStatistics()) {
            if (st.groupName == "scaling") {
                res->scalings_attempted += st.totalMoves;
                res->scalings_accepted += st.acceptedMoves;
            } else {
                res->moves_attempted += st.totalMoves;
                res->moves_accepted += st.acceptedMoves;
            }
        }
        res->density_mean = 0;
        res->density_error = 0;
        if (avgCycles > 0) {
            auto vals = sim->getObservablesCollector().getFlattenedAverageValues();
            if (!vals.empty()) {
                res->density_mean = vals.front().quantity.value;
                res->density_error = vals.front().quantity.error;
            }
        }
        const Packing &pk = sim->getPacking();
        res->ng_rebuild_seconds = pk.getNeighbourGridRebuildMicroseconds() * 1e-6;
        res->cycles = sim->getPerformedCycles();
        res->threads = threads;
        for (int i = 0; i < 3; i++) res->domains[i] = domainDivisions[i];

        const auto &dim = pk.getBox().getDimensions();
        if (boxOut)
            for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) boxOut[3*r + c] = dim(r, c);
        if (shapesOut)
            for (std::size_t i = 0; i < pk.size(); i++)
                fromShape(pk[i], shapesOut + 12*i);
    } catch (const std::exception &e) {
        if (err && errLen > 0) {
            std::strncpy(err, e.what(), errLen - 1);
            err[errLen - 1] = 0;
        }
        return 1;
    }
    return 0;
}

int rpr_omp_max_threads() { return OMP_MAXTHREADS; }

// `count` successive box proposals of one of the reference's scaler classes from a fresh std::mt19937(seed), every one
// from the same oldBox (core/volume_scalers/*.cpp, TriclinicBoxScaler::updateBox).
// kind: 1 delta_v, 2 linear, 3 log (AnisotropicVolumeScaler over the direction groups, bit masks 1 = x, 2 = y, 4 = z),
// 4 delta_triclinic(step, scaleTogether = !independent)
int rpr_propose_boxes(int kind, double step, int independent, const int *groups, int nGroups, std::uint64_t seed,
                      const double oldBox[9], int count, double *out, char *err, std::size_t errLen)
{
    try {
        std::unique_ptr<TriclinicBoxScaler> scaler;
        if (kind == 1) {
            scaler = std::make_unique<TriclinicAdapter>(std::make_unique<DeltaVolumeScaler>(), step);
        } else if (kind == 2 || kind == 3) {
            AnisotropicVolumeScaler::ScalingDirection dir;
            for (int g = 0; g < nGroups; g++) {
                AnisotropicVolumeScaler::ScalingDirection one;
                if (groups[g] & 1) one &= AnisotropicVolumeScaler::X;
                if (groups[g] & 2) one &= AnisotropicVolumeScaler::Y;
                if (groups[g] & 4) one &= AnisotropicVolumeScaler::Z;
                dir |= one;
            }
            std::unique_ptr<ScalingFactorSampler> sampler;
            if (kind == 2) sampler = std::make_unique<LinearScalingFactorSampler>();
            else sampler = std::make_unique<LogScalingFactorSampler>();
            scaler = std::make_unique<TriclinicAdapter>(
                std::make_unique<AnisotropicVolumeScaler>(std::move(sampler), dir, independent != 0), step);
        } else if (kind == 4) {
            scaler = std::make_unique<TriclinicDeltaScaler>(step, independent == 0);
        } else {
            throw std::runtime_error("unknown scaler kind");
        }
        std::mt19937 mt(seed);
        TriclinicBox box(mat(oldBox));
        for (int k = 0; k < count; k++) {
            TriclinicBox nb = scaler->updateBox(box, mt);
            const auto &dim = nb.getDimensions();
            for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) out[9*k + 3*r + c] = dim(r, c);
        }
    } catch (const std::exception &e) {
        if (err && errLen > 0) { std::strncpy(err, e.what(), errLen - 1); err[errLen - 1] = 0; }
        return 1;
    }
    return 0;
}

// Lattice(UnitCell(cellBox, molecules), dims) populated by SerialPopulator(axisOrder) with n shapes
// (core/lattice/Lattice.cpp:43-81, SerialPopulator.cpp:9-40); molecules: m x 12 (relative position, orientation)
int rpr_populate_lattice(const double cellBox[9], const double *molecules, std::uint64_t m, const int dims[3],
                         const char *axisOrder, std::uint64_t n, double boxOut[9], double *shapesOut, char *err,
                         std::size_t errLen)
{
    try {
        UnitCell cell(TriclinicBox(mat(cellBox)), toShapes(molecules, m));
        Lattice lattice(cell, {std::size_t(dims[0]), std::size_t(dims[1]), std::size_t(dims[2])});
        SerialPopulator populator{std::string(axisOrder)};
        auto shapes = populator.populateLattice(lattice, n);
        const TriclinicBox latticeBox = lattice.getLatticeBox();
        const auto &dim = latticeBox.getDimensions();
        for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) boxOut[3*r + c] = dim(r, c);
        for (std::size_t i = 0; i < shapes.size(); i++) fromShape(shapes[i], shapesOut + 12*i);
    } catch (const std::exception &e) {
        if (err && errLen > 0) { std::strncpy(err, e.what(), errLen - 1); err[errLen - 1] = 0; }
        return 1;
    }
    return 0;
}

}
