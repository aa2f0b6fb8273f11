// host/facade_check.cpp — TEST PROGRAM.  The same source is linked twice (host/Makefile): against the reference's own
// src/core/Packing.cpp (`facade_check_cpu`) and against host/Packing_b200.cpp + librampack_b200.so
// (`facade_check_b200`).  It drives a `Packing` through the reference's public interface only — the scenarios follow
// test/unit_tests/core/PackingTest.cpp (construction, explicit moves and accepts, scaling and revert, overlap
// counting, walls, energies, renormalisation, store / restore) with the reference's real ShapeTraits instead of its
// mock interactions — and prints one line per observed value.  tests/test_gpu_facade.py demands equal outputs:
// exact for verdicts, counts and stored positions, 1e-10 relative for energies.
#include <cstdio>
#include <iomanip>
#include <iostream>
#include <random>
#include <sstream>

#include "core/Packing.h"
#include "core/PeriodicBoundaryConditions.h"
#include "core/shapes/SphereTraits.h"
#include "core/shapes/SpherocylinderTraits.h"
#include "core/shapes/KMerTraits.h"
#include "core/interactions/RepulsiveLennardJonesInteraction.h"
#include "core/interactions/LennardJonesInteraction.h"

namespace {

void out(const std::string &key, double value) { std::printf("%s %.17g\n", key.c_str(), value); }

std::vector<Shape> lattice(const std::array<std::size_t, 3> &n, const std::array<double, 3> &cell, std::mt19937 &mt,
                           bool randomOrientations) {
    std::vector<Shape> shapes;
    std::uniform_real_distribution<double> unit(-1, 1);
    for (std::size_t i{}; i < n[0]; i++) for (std::size_t j{}; j < n[1]; j++) for (std::size_t k{}; k < n[2]; k++) {
        Vector<3> pos{(i + 0.5) * cell[0], (j + 0.5) * cell[1], (k + 0.5) * cell[2]};
        Matrix<3, 3> rot = Matrix<3, 3>::identity();
        if (randomOrientations) {
            Vector<3> axis;
            do { axis = {unit(mt), unit(mt), unit(mt)}; } while (axis.norm2() > 1 || axis.norm2() == 0);
            rot = Matrix<3, 3>::rotation(axis.normalized(), M_PI * unit(mt));
        }
        shapes.emplace_back(pos, rot);
    }
    return shapes;
}

void dumpState(const std::string &tag, const Packing &packing) {
    double sum{}, sumO{};
    std::size_t i{};
    for (const auto &shape : packing) {
        i++;
        for (int c = 0; c < 3; c++) sum += shape.getPosition()[c] * static_cast<double>(i % 7 + 1);
        for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) sumO += shape.getOrientation()(r, c) * (r + 2 * c + 1);
    }
    out(tag + ".size", static_cast<double>(packing.size()));
    out(tag + ".position_checksum", sum);
    out(tag + ".orientation_checksum", sumO);
    out(tag + ".first_x", packing.front().getPosition()[0]);
    out(tag + ".last_z", packing.back().getPosition()[2]);
    out(tag + ".volume", packing.getVolume());
}

void moveSequence(const std::string &tag, Packing &packing, const Interaction &interaction, std::mt19937 &mt,
                  std::size_t numMoves, double step, bool soft) {
    std::uniform_real_distribution<double> unit(-1, 1), u01(0, 1);
    std::uniform_int_distribution<std::size_t> pick(0, packing.size() - 1);
    std::size_t accepted{};
    double sumDE{};
    for (std::size_t m{}; m < numMoves; m++) {
        std::size_t idx = pick(mt);
        Vector<3> tr{step * unit(mt), step * unit(mt), step * unit(mt)};
        Vector<3> axis;
        do { axis = {unit(mt), unit(mt), unit(mt)}; } while (axis.norm2() > 1 || axis.norm2() == 0);
        Matrix<3, 3> rot = Matrix<3, 3>::rotation(axis.normalized(), 0.3 * unit(mt));
        double dE{};
        switch (m % 3) {
            case 0: dE = packing.tryTranslation(idx, tr, interaction); break;
            case 1: dE = packing.tryRotation(idx, rot, interaction); break;
            default: dE = packing.tryMove(idx, tr, rot, interaction); break;
        }
        double u = u01(mt);
        bool accept = u <= std::exp(-dE);
        if (std::isfinite(dE)) sumDE += dE;
        if (!soft) std::printf("%s.move%zu %s\n", tag.c_str(), m, std::isinf(dE) ? (dE > 0 ? "+inf" : "-inf") : "0");
        if (accept) {
            switch (m % 3) {
                case 0: packing.acceptTranslation(); break;
                case 1: packing.acceptRotation(); break;
                default: packing.acceptMove(); break;
            }
            accepted++;
        }
    }
    out(tag + ".accepted", static_cast<double>(accepted));
    out(tag + ".sum_finite_dE", sumDE);
}

void spherocylinders() {
    std::mt19937 mt(1234);
    SpherocylinderTraits traits(3, 0.5);
    const Interaction &interaction = traits.getInteraction();
    auto shapes = lattice({6, 6, 5}, {1.6, 1.6, 5.2}, mt, false);
    Packing packing({6 * 1.6, 6 * 1.6, 5 * 5.2}, shapes, std::make_unique<PeriodicBoundaryConditions>(), interaction);
    out("sc.overlaps0", static_cast<double>(packing.countTotalOverlaps(interaction, false)));
    moveSequence("sc.seq", packing, interaction, mt, 300, 0.4, false);
    out("sc.overlaps1", static_cast<double>(packing.countTotalOverlaps(interaction, false)));
    dumpState("sc.after_moves", packing);

    // scaling: down (overlaps -> +inf, reverted), up (accepted)
    double dE = packing.tryScaling(0.8, interaction);
    std::printf("sc.scale_down %s\n", std::isinf(dE) ? "+inf" : "0");
    packing.revertScaling();
    dumpState("sc.after_revert", packing);
    dE = packing.tryScaling(std::array<double, 3>{1.05, 1.02, 1.1}, interaction);
    out("sc.scale_up", dE);
    dumpState("sc.after_scale_up", packing);
    TriclinicBox sheared(Matrix<3, 3>{10.5, 0.4, 0, 0, 10.1, 0.3, 0.2, 0, 29.5});
    dE = packing.tryScaling(sheared, interaction);
    std::printf("sc.sheared %s\n", std::isinf(dE) ? "+inf" : "0");
    if (dE > 0) packing.revertScaling();
    dumpState("sc.after_shear", packing);

    // overlap counting (overlap relaxation): compress with overlaps allowed, then move
    packing.toggleOverlapCounting(true, interaction);
    out("sc.cached0", static_cast<double>(packing.getCachedNumberOfOverlaps()));
    dE = packing.tryScaling(0.85, interaction);
    std::printf("sc.counting_scale %s\n", std::isinf(dE) ? (dE > 0 ? "+inf" : "-inf") : "0");
    out("sc.cached1", static_cast<double>(packing.getCachedNumberOfOverlaps()));
    moveSequence("sc.counting_seq", packing, interaction, mt, 200, 0.6, false);
    out("sc.cached2", static_cast<double>(packing.getCachedNumberOfOverlaps()));
    out("sc.recount2", static_cast<double>(packing.countTotalOverlaps(interaction, false)));
    packing.toggleOverlapCounting(false, interaction);

    // walls
    packing.toggleWall(2, true);
    out("sc.wall_overlaps", static_cast<double>(packing.countWallOverlaps(interaction, false)));
    out("sc.total_with_walls", static_cast<double>(packing.countTotalOverlaps(interaction, false)));
    packing.toggleWall(2, false);

    // store / restore round trip
    std::stringstream stream;
    packing.store(stream, {{"cycles", "17"}});
    {   // RAMSNAP text (io/RamsnapWriter.cpp:8-48): byte-identical files for identical state
        std::uint64_t hash = 1469598103934665603ull;
        for (unsigned char ch : stream.str()) hash = (hash ^ ch) * 1099511628211ull;
        std::printf("sc.ramsnap_bytes %zu\nsc.ramsnap_fnv1a %016llx\n", stream.str().size(),
                    static_cast<unsigned long long>(hash));
    }
    Packing restored(std::make_unique<PeriodicBoundaryConditions>());
    auto aux = restored.restore(stream, interaction);
    std::printf("sc.aux_cycles %s\n", aux["cycles"].c_str());
    dumpState("sc.restored", restored);
    out("sc.neighbours", packing.getAverageNumberOfNeighbours());
}

void spheresRenormalise() {
    std::mt19937 mt(99);
    SphereTraits traits(0.5);
    const Interaction &interaction = traits.getInteraction();
    auto shapes = lattice({7, 7, 7}, {1.3, 1.3, 1.3}, mt, true);
    Packing packing({7 * 1.3, 7 * 1.3, 7 * 1.3}, shapes, std::make_unique<PeriodicBoundaryConditions>(), interaction);
    moveSequence("sph.seq", packing, interaction, mt, 300, 0.3, false);
    out("sph.rejected_fixes", static_cast<double>(packing.renormalizeOrientations(interaction, false)));
    dumpState("sph.after_fix", packing);
    out("sph.packing_fraction", packing.getPackingFraction(traits.getGeometry().getVolume()));
}

void softTrimers() {
    std::mt19937 mt(7);
    KMerTraits traits(3, 0.5, 1, std::make_shared<RepulsiveLennardJonesInteraction>(1, 1));
    const Interaction &interaction = traits.getInteraction();
    auto shapes = lattice({5, 5, 5}, {1.25, 1.25, 3.2}, mt, false);
    Packing packing({5 * 1.25, 5 * 1.25, 5 * 3.2}, shapes, std::make_unique<PeriodicBoundaryConditions>(), interaction);
    out("wca.energy0", packing.getTotalEnergy(interaction));
    moveSequence("wca.seq", packing, interaction, mt, 300, 0.2, true);
    out("wca.energy1", packing.getTotalEnergy(interaction));
    out("wca.fluctuations", packing.getParticleEnergyFluctuations(interaction));
    double dE = packing.tryScaling(0.97, interaction);
    out("wca.scale_dE", dE);
    packing.revertScaling();
    out("wca.energy2", packing.getTotalEnergy(interaction));
    dumpState("wca.final", packing);
}

}  // namespace

int main() {
    try {
        spherocylinders();
        spheresRenormalise();
        softTrimers();
    } catch (const std::exception &ex) {
        std::printf("EXCEPTION %s\n", ex.what());
        return 1;
    }
    return 0;
}
