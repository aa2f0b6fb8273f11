// rp_sim.cu — the host cycle loop in C++ (no kernels here): Simulation::performCycle, tryScaling and
// evaluateCounters of the reference (core/Simulation.cpp:325-363, 493-531, 533-607) and its box scalers
// (core/volume_scalers/*.cpp), driving rp_sweep / rp_try_scaling of the same handle.  One box proposal and one
// Metropolis test per cycle run on the host; everything per-particle runs on the device.
#include "rp_internal.cuh"
#include <cmath>
#include <cstring>
#include <random>

namespace rp {
Box makeBox(const double m[9]);      // rp_api.cu
extern thread_local std::string tlsError;
}

struct rp_sim {
    rp_handle *h = nullptr;
    rp_sim_config cfg;
    std::mt19937 mt;
    rp_sim_stats st;
    // Simulation::Counter (core/Simulation.h:60-86): moves / accepted since the last evaluation
    unsigned long long curMoves[RP_SIM_MAX_SAMPLERS] = {}, curAccepted[RP_SIM_MAX_SAMPLERS] = {};
    unsigned long long curScalings = 0, curScalingsAccepted = 0;
    std::string error;
    // Sweeps of the specialised kernel are queued without a host synchronisation: their counters are accumulated on
    // the device ({attempted, accepted, left the cell, overlap} per sampler) and read when somebody needs them —
    // evaluateCounters once a window of 100 N moves is full, rp_sim_stats_get, the end of rp_sim_run.
    unsigned long long *dAcc = nullptr;                         // 4 * RP_SIM_MAX_SAMPLERS
    unsigned long long devSeen[4 * RP_SIM_MAX_SAMPLERS] = {};   // device accumulators at the last read
    unsigned long long pendingMoves[RP_SIM_MAX_SAMPLERS] = {};  // attempted moves queued since the last read (estimate)
    double accSweepBase = 0, accGridBase = 0;                   // h->accSweepMs / accSweepGridMs at creation
    int device = 0;                                             // (the handle may be gone when the rp_sim is destroyed)
};

namespace {
using namespace rp;

constexpr double STEP_MODIFY_FACTOR = 1.1;      // TriclinicBoxScaler::STEP_MODIFY_FACTOR, MoveSampler step changes

// TriclinicBox(sides) with every side multiplied by its factor (TriclinicAdapter.h:27-33)
void scaleSides(const double oldBox[9], const double f[3], double out[9]) {
    for (int r = 0; r < 3; r++)
        for (int c = 0; c < 3; c++) out[3 * r + c] = oldBox[3 * r + c] * f[c];
}

void sampleLinearOrLog(rp_sim *s, bool logarithmic, int mask, const double heights[3], double f[3]) {
    std::uniform_real_distribution<double> dist(-s->st.scaler_step, s->st.scaler_step);
    const double x = dist(s->mt);
    for (int i = 0; i < 3; i++) {
        f[i] = 1;
        if (!((mask >> i) & 1)) continue;
        f[i] = logarithmic ? std::exp(x) : (heights[i] + x) / heights[i];
        if (!(f[i] > 0)) throw rp_error(RP_ERR_PRECONDITION, "Ensures(scalingFactors[i] > 0)");
    }
}

void proposeBox(rp_sim *s, const double oldBox[9], double out[9]) {
    const rp_sim_config &c = s->cfg;
    Box b = makeBox(oldBox);
    double heights[3];
    boxHeights(b, heights);
    switch (c.scaler_kind) {
        case RP_SCALER_DELTA_V: {
            // DeltaVolumeScaler.cpp:10-19 — "currentV" is the product of the heights handed over by TriclinicAdapter
            std::uniform_real_distribution<double> unit(0, 1);
            double deltaV = (2 * unit(s->mt) - 1) * s->st.scaler_step;
            double currentV = heights[0] * heights[1] * heights[2];
            double factor = std::cbrt((deltaV + currentV) / currentV);
            double f[3] = {factor, factor, factor};
            scaleSides(oldBox, f, out);
            return;
        }
        case RP_SCALER_LINEAR:
        case RP_SCALER_LOG: {
            const bool logarithmic = c.scaler_kind == RP_SCALER_LOG;
            double f[3] = {1, 1, 1};
            if (c.independent && c.n_direction_groups > 1) {
                std::uniform_int_distribution<std::size_t> pick(0, c.n_direction_groups - 1);
                sampleLinearOrLog(s, logarithmic, c.direction_groups[pick(s->mt)], heights, f);
            } else {
                for (int g = 0; g < c.n_direction_groups; g++) {
                    double single[3];
                    sampleLinearOrLog(s, logarithmic, c.direction_groups[g], heights, single);
                    for (int i = 0; i < 3; i++) f[i] *= single[i];
                }
            }
            scaleSides(oldBox, f, out);
            return;
        }
        case RP_SCALER_TRICLINIC_DELTA: {
            // TriclinicDeltaScaler.cpp:7-25
            std::memcpy(out, oldBox, 9 * sizeof(double));
            std::uniform_real_distribution<double> step(-s->st.scaler_step, s->st.scaler_step);
            if (!c.independent) {
                for (int i = 0; i < 3; i++)
                    for (int j = 0; j < 3; j++) out[3 * i + j] += step(s->mt);
            } else {
                std::uniform_int_distribution<std::size_t> sideDist(0, 2);
                std::size_t side = sideDist(s->mt);
                for (int i = 0; i < 3; i++) out[3 * i + side] += step(s->mt);
            }
            return;
        }
        default:
            throw rp_error(RP_ERR_PRECONDITION, "box scaling is disabled for this rp_sim");
    }
}

void check(rp_sim *s, rp_status st) {
    if (st != RP_OK) throw rp_error(st, rp_last_error(s->h));
}

// Simulation::tryScaling (core/Simulation.cpp:493-531)
bool tryScalingStep(rp_sim *s) {
    rp_handle *h = s->h;
    double oldBox[9], newBox[9];
    for (int i = 0; i < 9; i++) oldBox[i] = h->box.m.a[i];
    proposeBox(s, oldBox, newBox);
    Box nb = makeBox(newBox);
    const double oldV = boxVolume(h->box), newV = boxVolume(nb);
    RP_EXPECT(newV != 0, "newBox.getVolume() != 0");
    const double deltaV = newV - oldV, factor = newV / oldV, N = static_cast<double>(h->n);
    const double T = s->cfg.temperature, p = s->cfg.pressure;
    std::uniform_real_distribution<double> unit(0, 1);
    const bool soft = h->inter.soft != RP_SOFT_NONE || h->overlapCounting;
    double dE = 0;
    if (!soft) {
        // hard-only shortcut: the volume term alone decides first, the overlap check only runs on survivors
        double exponent = N * std::log(factor) - p * deltaV / T;
        if (unit(s->mt) > std::exp(exponent)) return false;
        check(s, rp_try_scaling(h, newBox, &dE));
        s->st.scalings_evaluated++;
        s->st.scaling_ms += h->lastScalingMs;
        if (dE > 0) { check(s, rp_revert_scaling(h)); return false; }
        return true;
    }
    check(s, rp_try_scaling(h, newBox, &dE));
    s->st.scalings_evaluated++;
    s->st.scaling_ms += h->lastScalingMs;
    double exponent = N * std::log(factor) - dE / T - p * deltaV / T;
    if (unit(s->mt) <= std::exp(exponent)) return true;
    check(s, rp_revert_scaling(h));
    return false;
}

// device accumulators -> host counters (blocks until the queued sweeps are done)
void readDeviceCounters(rp_sim *s) {
    bool any = false;
    for (int i = 0; i < s->cfg.n_samplers; i++) any = any || s->pendingMoves[i] > 0;
    if (!any) return;
    unsigned long long now[4 * RP_SIM_MAX_SAMPLERS];
    readbackSync(s->h, s->h->stream, now, s->dAcc, sizeof(now));
    RP_CUDA(cudaGetLastError());
    for (int i = 0; i < s->cfg.n_samplers; i++) {
        const unsigned long long att = now[4 * i] - s->devSeen[4 * i], acc = now[4 * i + 1] - s->devSeen[4 * i + 1];
        s->st.moves_attempted[i] += att;
        s->st.moves_accepted[i] += acc;
        s->curMoves[i] += att;
        s->curAccepted[i] += acc;
        s->pendingMoves[i] = 0;
    }
    std::memcpy(s->devSeen, now, sizeof(now));
}

// Simulation::evaluateMoleculeMoveCounter / evaluateScalingMoveCounter (core/Simulation.cpp:533-607)
void evaluateCounters(rp_sim *s) {
    const unsigned long long N = s->h->n;
    for (int i = 0; i < s->cfg.n_samplers; i++)
        if (s->curMoves[i] + s->pendingMoves[i] >= 100 * N) { readDeviceCounters(s); break; }
    double heights[3];
    boxHeights(s->h->box, heights);
    const double maxTranslation = std::min(heights[0], std::min(heights[1], heights[2])) / 2;
    for (int i = 0; i < s->cfg.n_samplers; i++) {
        if (s->cfg.samplers[i].kind == RP_MOVE_FLIP) continue;      // FlipSampler::increase/decreaseStepSize: no-ops
        if (s->curMoves[i] < 100 * N) continue;
        const double rate = static_cast<double>(s->curAccepted[i]) / static_cast<double>(s->curMoves[i]);
        s->curMoves[i] = s->curAccepted[i] = 0;
        const int kind = s->cfg.samplers[i].kind;
        double &ts = s->st.translation_step[i], &rs = s->st.rotation_step[i];
        if (rate > 0.2) {
            if (kind == RP_MOVE_ROTATION) { if (rs < M_PI) rs *= STEP_MODIFY_FACTOR; }     // RotationSampler.cpp
            else if (!(ts >= maxTranslation)) {       // (Roto)TranslationSampler::increaseTranslationStepSize
                ts *= STEP_MODIFY_FACTOR;
                if (kind == RP_MOVE_ROTOTRANSLATION) rs *= STEP_MODIFY_FACTOR;
            }
        } else if (rate < 0.1) {
            if (kind != RP_MOVE_ROTATION) ts /= STEP_MODIFY_FACTOR;
            if (kind != RP_MOVE_TRANSLATION) rs /= STEP_MODIFY_FACTOR;
        }
    }
    if (s->cfg.scaler_kind != RP_SCALER_NONE && s->curScalings >= 100) {
        const double rate = static_cast<double>(s->curScalingsAccepted) / static_cast<double>(s->curScalings);
        s->curScalings = s->curScalingsAccepted = 0;
        if (rate > 0.2) s->st.scaler_step *= STEP_MODIFY_FACTOR;
        else if (rate < 0.1) s->st.scaler_step /= STEP_MODIFY_FACTOR;
    }
}

void performCycle(rp_sim *s, bool adjust) {
    rp_handle *h = s->h;
    double heights[3];
    boxHeights(h->box, heights);
    const double maxTranslation = std::min(heights[0], std::min(heights[1], heights[2])) / 2;
    for (int i = 0; i < s->cfg.n_samplers; i++) {
        rp_sweep_params p;
        std::memset(&p, 0, sizeof(p));
        p.kind = s->cfg.samplers[i].kind;
        p.pass = i;
        // MoveSampler step clamps (TranslationSampler.cpp:20-37, RototranslationSampler.cpp:33)
        p.translation_step = std::min(s->st.translation_step[i], maxTranslation);
        p.rotation_step = std::min(s->st.rotation_step[i], M_PI);
        p.temperature = s->cfg.temperature;
        // the Philox streams of a sweep are keyed by (handle seed, particle, cycle, pass): the cycle number is the
        // handle's own count of cycles, so a second rp_sim on the same handle (thermalisation, then averaging)
        // continues the streams instead of replaying them
        p.cycle = h->simCycles;
        for (int q = 0; q < 3; q++) {
            p.flip_axis[q] = s->cfg.samplers[i].flip_axis[q];
            p.flip_origin[q] = s->cfg.samplers[i].flip_origin[q];
        }
        p.flip_every = s->cfg.samplers[i].flip_every;
        if (sweepIsQueued(h, p.kind)) {
            sweepAsync(h, p, s->dAcc + 4 * i, nullptr);
            s->pendingMoves[i] += h->n;
        } else {
            rp_sweep_counters c;
            sweepAsync(h, p, nullptr, &c);
            s->st.moves_attempted[i] += c.attempted;
            s->st.moves_accepted[i] += c.accepted;
            s->curMoves[i] += c.attempted;
            s->curAccepted[i] += c.accepted;
        }
    }
    if (s->cfg.scaler_kind != RP_SCALER_NONE) {
        bool wasScaled = tryScalingStep(s);
        s->st.scalings_attempted++;
        s->curScalings++;
        if (wasScaled) { s->st.scalings_accepted++; s->curScalingsAccepted++; }
    }
    if (adjust) evaluateCounters(s);
    s->st.cycles++;
    h->simCycles++;
    // Simulation::fixRotationMatrices (core/Simulation.cpp:144-145, 750-769)
    if (s->cfg.rotation_matrix_fix_every > 0 && s->st.cycles % s->cfg.rotation_matrix_fix_every == 0)
        check(s, rp_renormalize_orientations(h, h->overlapCounting ? 1 : 0, nullptr));
}
}  // namespace

extern "C" {

rp_status rp_sim_create(rp_handle *h, const rp_sim_config *config, rp_sim **out) {
    if (!config || !out) { rp::tlsError = "rp_sim_create: config and out must not be NULL"; return RP_ERR_PRECONDITION; }
    const rp_sim_config &c = *config;
    bool ok = c.n_samplers >= 0 && c.n_samplers <= RP_SIM_MAX_SAMPLERS && c.temperature > 0
              && c.scaler_kind >= RP_SCALER_NONE && c.scaler_kind <= RP_SCALER_TRICLINIC_DELTA
              && (c.scaler_kind == RP_SCALER_NONE || c.scaler_step > 0);
    if (c.scaler_kind == RP_SCALER_LINEAR || c.scaler_kind == RP_SCALER_LOG) {
        ok = ok && c.n_direction_groups >= 1 && c.n_direction_groups <= 3;
        for (int g = 0; ok && g < c.n_direction_groups; g++) ok = c.direction_groups[g] >= 1 && c.direction_groups[g] <= 7;
    }
    for (int i = 0; ok && i < c.n_samplers; i++) {
        int k = c.samplers[i].kind;
        ok = k >= RP_MOVE_TRANSLATION && k <= RP_MOVE_FLIP
             && (k == RP_MOVE_ROTATION || k == RP_MOVE_FLIP || c.samplers[i].translation_step > 0)
             && (k == RP_MOVE_TRANSLATION || k == RP_MOVE_FLIP || c.samplers[i].rotation_step > 0)
             && (k != RP_MOVE_FLIP || c.samplers[i].flip_every > 0);
    }
    if (!ok) { rp::tlsError = "rp_sim_create: invalid configuration (steps > 0, temperature > 0, known kinds)"; return RP_ERR_PRECONDITION; }
    rp_sim *s = new rp_sim;
    s->h = h;
    s->cfg = c;
    s->mt.seed(static_cast<std::mt19937::result_type>(c.seed));
    std::memset(&s->st, 0, sizeof(s->st));
    for (int i = 0; i < c.n_samplers; i++) {
        s->st.translation_step[i] = c.samplers[i].translation_step;
        s->st.rotation_step[i] = c.samplers[i].rotation_step;
    }
    s->st.scaler_step = c.scaler_step;
    if (h) {
        cudaSetDevice(h->device);
        bool ok = cudaMalloc(&s->dAcc, 4 * RP_SIM_MAX_SAMPLERS * sizeof(unsigned long long)) == cudaSuccess;
        if (ok) {
            try {
                zeroAsync(h, s->dAcc, 4 * RP_SIM_MAX_SAMPLERS * sizeof(unsigned long long), h->stream);
            } catch (const std::exception &) {
                ok = false;
            }
        }
        if (!ok) {
            rp::tlsError = "rp_sim_create: cannot allocate the device counters";
            delete s;
            return RP_ERR_CUDA;
        }
        s->device = h->device;
        s->accSweepBase = h->accSweepMs;
        s->accGridBase = h->accSweepGridMs;
    }
    *out = s;
    return RP_OK;
}

void rp_sim_destroy(rp_sim *s) {
    if (!s) return;
    if (s->dAcc) {
        // no access to the handle here: callers (garbage collectors) may destroy it first.  cudaFree waits for the
        // device work that may still touch the buffer.
        cudaSetDevice(s->device);
        cudaFree(s->dAcc);
    }
    delete s;
}

rp_status rp_sim_run(rp_sim *s, uint64_t cycles, int adjust_step_sizes) {
    if (!s || !s->h) { rp::tlsError = "rp_sim_run: needs an rp_sim bound to a handle"; return RP_ERR_PRECONDITION; }
    try {
        RP_EXPECT(s->h->hasInteraction && s->h->n > 0, "packing and interaction must be set");
        for (uint64_t c = 0; c < cycles; c++) performCycle(s, adjust_step_sizes != 0);
        // one synchronisation per call: counters, device times, staging-capacity reports of the queued sweeps
        readDeviceCounters(s);
        RP_CUDA(cudaStreamSynchronize(s->h->stream));
        sweepTimings(s->h, true);
        sweepHarvest(s->h, true);
    } catch (const rp::rp_error &e) {
        s->h->error = e.what();
        return static_cast<rp_status>(e.status);
    } catch (const std::exception &e) {
        s->h->error = e.what();
        return RP_ERR_STATE;
    }
    return RP_OK;
}

rp_status rp_sim_stats_get(const rp_sim *s, rp_sim_stats *out) {
    if (!s || !out) return RP_ERR_PRECONDITION;
    *out = s->st;
    if (s->h) {
        for (int i = 0; i < 9; i++) out->box[i] = s->h->box.m.a[i];
        out->sweep_kernel_ms = s->h->accSweepMs - s->accSweepBase;
        out->sweep_grid_ms = s->h->accSweepGridMs - s->accGridBase;
        out->cells_skipped = s->h->sweepCellsSkipped;
    }
    return RP_OK;
}

rp_status rp_sim_propose_box(rp_sim *s, const double old_box[9], double new_box[9]) {
    if (!s || !old_box || !new_box) return RP_ERR_PRECONDITION;
    try {
        proposeBox(s, old_box, new_box);
    } catch (const rp::rp_error &e) {
        rp::tlsError = e.what();
        return static_cast<rp_status>(e.status);
    }
    return RP_OK;
}

}  // extern "C"
