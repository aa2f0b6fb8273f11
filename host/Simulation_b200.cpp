// host/Simulation_b200.cpp — the three link-time hooks that turn the reference's per-move cycle into bulk device sweeps.
//
// The reference's src/core/Simulation.cpp is compiled UNMODIFIED; host/Makefile then post-processes its object file:
//   * Simulation::performCycle is made a weak symbol, so the definition below replaces it for every caller
//     (integrate, relaxOverlaps — src/core/Simulation.cpp:139, 169, 256);
//   * Simulation::integrate / relaxOverlaps (the Environment + parameter-struct overloads the front end calls,
//     src/frontend/modes/CasinoMode.cpp:223, 297) are renamed to rampack_ref_Simulation_*; the wrappers below note
//     the cadence at which observers read the packing, call the reference body and bring the host mirror up to date.
// Everything else of Simulation — tryScaling and the NpT rule, evaluateCounters and the step adaptation, thermodynamic
// parameter updates, observers, recorders, inline info, signal handling — runs from the reference's own object code.
//
// Compiled with -fno-access-control (it defines and uses private members of Simulation and reads sampler parameters).
#include <algorithm>
#include <chrono>
#include <cmath>
#include <map>

#include "core/Simulation.h"
#include "core/move_samplers/TranslationSampler.h"
#include "core/move_samplers/RotationSampler.h"
#include "core/move_samplers/RototranslationSampler.h"
#include "core/move_samplers/FlipSampler.h"
#include "utils/Exceptions.h"

#include "rampack_b200.h"
#include "packing_b200.h"

// the reference bodies under their new names (host/Makefile: objcopy --redefine-sym).  A non-static member function
// and a free function taking the object pointer first share one calling convention in the Itanium C++ ABI.
void rampack_ref_Simulation_integrate(Simulation *self, Simulation::Environment env,
                                      const Simulation::IntegrationParameters &params, const ShapeTraits &shapeTraits,
                                      std::shared_ptr<ObservablesCollector> observablesCollector,
                                      std::vector<std::unique_ptr<SimulationRecorder>> simulationRecorders,
                                      Logger &logger) asm("rampack_ref_Simulation_integrate");
void rampack_ref_Simulation_relaxOverlaps(Simulation *self, Simulation::Environment env,
                                          const Simulation::OverlapRelaxationParameters &params,
                                          const ShapeTraits &shapeTraits,
                                          std::shared_ptr<ObservablesCollector> observablesCollector,
                                          std::vector<std::unique_ptr<SimulationRecorder>> simulationRecorders,
                                          Logger &logger) asm("rampack_ref_Simulation_relaxOverlaps");

namespace {

// when does somebody read Packing::shapes between two cycles?  (snapshots, averaging, inline info, matrix fix)
struct Cadence {
    bool known{};
    std::vector<std::size_t> every;
};
std::map<const Simulation *, Cadence> cadences;
std::map<const Simulation *, std::uint64_t> sweepCycles;      // Philox cycle counter: never restarts within a process

struct CadenceScope {
    const Simulation *sim;
    CadenceScope(const Simulation *sim, std::vector<std::size_t> every) : sim{sim} {
        Cadence c;
        c.known = true;
        for (auto e : every) if (e > 0) c.every.push_back(e);
        cadences[sim] = c;
    }
    ~CadenceScope() { cadences.erase(sim); }
};

bool describeSampler(MoveSampler &sampler, const Packing &packing, rp_sweep_params &p) {
    auto heights = packing.getBox().getHeights();
    double maxTranslation = *std::min_element(heights.begin(), heights.end()) / 2;
    if (auto *t = dynamic_cast<TranslationSampler *>(&sampler)) {
        // TranslationSampler.cpp:20-37: step = min(maxTranslationStepSize, translationStepSize), max implicit = h/2
        p.kind = RP_MOVE_TRANSLATION;
        double maxStep = t->isMaxTranslationStepSizeImplicit ? maxTranslation : t->maxTranslationStepSize;
        t->maxTranslationStepSize = maxStep;
        p.translation_step = std::min(maxStep, t->translationStepSize);
        return true;
    }
    if (auto *r = dynamic_cast<RotationSampler *>(&sampler)) {
        p.kind = RP_MOVE_ROTATION;
        p.rotation_step = std::min(r->rotationStepSize, M_PI);
        return true;
    }
    if (auto *rt = dynamic_cast<RototranslationSampler *>(&sampler)) {
        // RototranslationSampler.cpp:22-49
        p.kind = RP_MOVE_ROTOTRANSLATION;
        double maxStep = rt->isMaxTranslationStepSizeImplicit ? maxTranslation : rt->maxTranslationStepSize;
        rt->maxTranslationStepSize = maxStep;
        p.translation_step = std::min(maxStep, rt->translationStepSize);
        p.rotation_step = std::min(*rt->rotationStepSize, M_PI);
        return true;
    }
    if (auto *f = dynamic_cast<FlipSampler *>(&sampler)) {
        // FlipSampler.cpp:18-51
        p.kind = RP_MOVE_FLIP;
        p.flip_every = f->flipEvery;
        for (int c = 0; c < 3; c++) {
            p.flip_axis[c] = f->flipAxis[c];
            p.flip_origin[c] = f->geometricOrigin[c];
        }
        return true;
    }
    return false;
}

}  // namespace

// Simulation::performMoves (src/core/Simulation.cpp:365-491) as one bulk device sweep per MoveSampler.  Returns false
// when the device cannot sweep this configuration (box lower than 2 x totalRange along an axis, unknown sampler):
// the caller then runs the reference's own per-move loop, which reaches the device move by move through Packing.
static bool performMovesOnDevice(Simulation &sim) {
    Packing &packing = *sim.packing;
    rp_handle *h = rampack_b200::handle_of(packing);
    auto &samplers = sim.environment.moveSamplers;
    std::uint64_t &cycle = sweepCycles[&sim];
    for (std::size_t i{}; i < samplers.size(); i++) {
        rp_sweep_params p{};
        if (!describeSampler(*samplers[i], packing, p)) return false;
        p.pass = static_cast<int32_t>(i);
        p.temperature = sim.temperature;
        p.cycle = cycle;
        rp_sweep_counters c{};
        rp_status status = rp_sweep(h, &p, &c);
        if (status == RP_ERR_UNSUPPORTED && i == 0) return false;
        rampack_b200::check(h, status);
        rampack_b200::mark_mirror_stale(packing);
        auto &counter = sim.moveCounters[i];
        counter.movesSinceEvaluation += c.attempted;
        counter.acceptedMovesSinceEvaluation += c.accepted;
        counter.moves += c.attempted;
        counter.acceptedMoves += c.accepted;
    }
    cycle++;
    return true;
}

void Simulation::performCycle(Logger &logger, const ShapeTraits &shapeTraits) {
    const auto &interaction = shapeTraits.getInteraction();
    // inside a cycle nobody reads the host mirror: the Packing calls below leave it stale
    struct LazyScope {
        const Packing &packing;
        bool old;
        explicit LazyScope(const Packing &packing) : packing{packing}, old{rampack_b200::set_lazy_mirror(packing, true)} { }
        ~LazyScope() { rampack_b200::set_lazy_mirror(packing, old); }
    } lazyScope(*this->packing);

    using namespace std::chrono;
    auto start = high_resolution_clock::now();
    if (!performMovesOnDevice(*this)) {
        // (the per-move path keeps the mirror current move by move: Packing::acceptMove)
        rampack_b200::sync_mirror(*this->packing);
        rampack_b200::set_lazy_mirror(*this->packing, false);
        this->performMovesWithoutDomainDivision(shapeTraits);
        rampack_b200::set_lazy_mirror(*this->packing, true);
    }
    auto end = high_resolution_clock::now();
    this->moveMicroseconds += duration<double, std::micro>(end - start).count();

    if (this->environment.isBoxScalingEnabled()) {
        start = high_resolution_clock::now();
        bool wasScaled = this->tryScaling(interaction);
        this->scalingCounter.increment(wasScaled);
        end = high_resolution_clock::now();
        this->scalingMicroseconds += duration<double, std::micro>(end - start).count();
    }

    if (this->shouldAdjustStepSize)
        this->evaluateCounters(logger);

    this->performedCycles++;
    this->totalCycles++;

    // The observers read Packing::shapes through the header's inline begin() / end(): the mirror is refreshed after
    // every cycle at whose end somebody may look (every cycle when the cadence is unknown)
    auto it = cadences.find(this);
    bool somebodyLooks = it == cadences.end() || !it->second.known || this->totalCycles == this->maxCycles;
    if (!somebodyLooks)
        for (auto every : it->second.every)
            somebodyLooks = somebodyLooks || this->totalCycles % every == 0;
    if (somebodyLooks)
        rampack_b200::sync_mirror(*this->packing);
}

void Simulation::integrate(Environment env, const IntegrationParameters &params, const ShapeTraits &shapeTraits,
                           std::shared_ptr<ObservablesCollector> observablesCollector_,
                           std::vector<std::unique_ptr<SimulationRecorder>> simulationRecorders, Logger &logger)
{
    CadenceScope scope(this, {params.snapshotEvery, params.averagingCycles > 0 ? params.averagingEvery : 0,
                              params.inlineInfoEvery, params.rotationMatrixFixEvery});
    rampack_ref_Simulation_integrate(this, std::move(env), params, shapeTraits, std::move(observablesCollector_),
                                     std::move(simulationRecorders), logger);
    rampack_b200::sync_mirror(*this->packing);
}

void Simulation::relaxOverlaps(Environment env, const OverlapRelaxationParameters &params,
                               const ShapeTraits &shapeTraits,
                               std::shared_ptr<ObservablesCollector> observablesCollector_,
                               std::vector<std::unique_ptr<SimulationRecorder>> simulationRecorders, Logger &logger)
{
    CadenceScope scope(this, {params.snapshotEvery, params.inlineInfoEvery, params.rotationMatrixFixEvery});
    rampack_ref_Simulation_relaxOverlaps(this, std::move(env), params, shapeTraits, std::move(observablesCollector_),
                                         std::move(simulationRecorders), logger);
    rampack_b200::sync_mirror(*this->packing);
}
