// host/Packing_b200.cpp — the reference's `Packing` (src/core/Packing.h, UNMODIFIED header) implemented over the C ABI
// of librampack_b200.so.  Linked INSTEAD of the reference's src/core/Packing.cpp underneath the reference's unmodified
// Simulation, observables, I/O and PYON front end (host/Makefile): every particle-level operation runs on the device,
// the object keeps a host mirror of the shapes for the read side (observers, recorders, RAMSNAP writer).
//
//   method here                                   reference body it replaces            C ABI
//   Packing::Packing / reset                      Packing.cpp:46-82                     rp_create, rp_set_interaction, rp_upload
//   setupForInteraction                           Packing.cpp:931-950                   rp_set_interaction
//   tryTranslation / tryRotation / tryMove        Packing.cpp:84-164                    rp_try_move
//   acceptTranslation / acceptRotation / Move     Packing.cpp:234-302                   rp_accept_move (+ rp_download_range)
//   tryScaling / revertScaling                    Packing.cpp:166-208, 363-371          rp_try_scaling / rp_revert_scaling
//   countTotalOverlaps / countWallOverlaps        Packing.cpp:519-580                   rp_count_total_overlaps / rp_count_wall_overlaps
//   getTotalEnergy / getParticleEnergyFluct.      Packing.cpp:690-709, 841-856          rp_total_energy / rp_particle_energy_fluctuations
//   toggleOverlapCounting / getCachedNumber...    Packing.cpp:1049-1060                 rp_toggle_overlap_counting / rp_cached_overlaps
//   toggleWall(s)                                 Packing.cpp:1075-1084                 rp_toggle_walls
//   renormalizeOrientations                       Packing.cpp:1178-1212                 rp_renormalize_orientations
//   getAverageNumberOfNeighbours                  Packing.cpp:1000-1033                 rp_neighbour_stats
//
// The device handle and the mirror state live in a BoundaryConditions decorator that the Packing owns through its
// `bc` member (the header has no spare member and no user-declared destructor; the decorator dies with the Packing).
// Errors: rp_status -> the reference's exception types (PreconditionException for Expects, std::runtime_error).
#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>
#include <sstream>

#include "core/Packing.h"
#include "core/io/RamsnapWriter.h"
#include "core/io/RamsnapReader.h"
#include "utils/Exceptions.h"
#include "utils/Utils.h"

#include "rampack_b200.h"
#include "describe_interaction.h"
#include "packing_b200.h"

static_assert(sizeof(Shape) == 12 * sizeof(double), "Shape = Vector<3> position + Matrix<3, 3> orientation (Shape.h:20-24)");

namespace rampack_b200 {

// BoundaryConditions decorator: forwards to the user's object, owns the device side of the Packing
class DeviceState : public BoundaryConditions {
public:
    std::unique_ptr<BoundaryConditions> inner;
    rp_handle *handle{};
    bool mirrorStale{};            // the device holds newer shapes than Packing::shapes
    // Readers reach Packing::shapes through the header's INLINE begin() / end(), which no facade code can intercept.
    // So by default every mutating Packing call refreshes the mirror before it returns (eager: the compatibility path
    // of unit tests and tools).  The Simulation hooks switch to lazy mode for the duration of a cycle and refresh the
    // mirror themselves whenever an observer may look (Simulation_b200.cpp).
    bool lazy{};
    bool uploaded{};
    rp_interaction_desc desc{};
    rp_xc_program xc{};
    bool hasDesc{};
    std::size_t lastAlteredKind{};

    explicit DeviceState(std::unique_ptr<BoundaryConditions> inner) : inner{std::move(inner)} {
        rp_config config{};
        config.device = 0;
        if (const char *dev = std::getenv("RAMPACK_B200_DEVICE")) config.device = std::atoi(dev);
        config.seed = 1234;
        if (const char *seed = std::getenv("RAMPACK_B200_SEED")) config.seed = std::strtoull(seed, nullptr, 10);
        if (rp_create(&config, &this->handle) != RP_OK)
            throw std::runtime_error(std::string("rampack_b200: ") + rp_last_error(nullptr));
    }
    ~DeviceState() override { rp_destroy(this->handle); }

    void setBox(const TriclinicBox &box) override { this->inner->setBox(box); }
    [[nodiscard]] Vector<3> getCorrection(const Vector<3> &position) const override {
        return this->inner->getCorrection(position);
    }
    [[nodiscard]] Vector<3> getTranslation(const Vector<3> &position1, const Vector<3> &position2) const override {
        return this->inner->getTranslation(position1, position2);
    }
    [[nodiscard]] double getDistance2(const Vector<3> &position1, const Vector<3> &position2) const override {
        return this->inner->getDistance2(position1, position2);
    }
};

void check(rp_handle *h, rp_status st) {
    if (st == RP_OK) return;
    std::string msg = rp_last_error(h);
    if (st == RP_ERR_PRECONDITION)
        throw PreconditionException("librampack_b200", "rp_*", 0, msg, "");     // what Expects() throws (utils/Exceptions.h:19-66)
    throw std::runtime_error("rampack_b200: " + msg);
}

DeviceState &state(const Packing &packing) {
    auto *st = dynamic_cast<DeviceState *>(const_cast<BoundaryConditions *>(&packing.getBoundaryConditions()));
    if (st == nullptr) throw std::runtime_error("rampack_b200: this Packing was not created by Packing_b200.cpp");
    return *st;
}

rp_handle *handle_of(const Packing &packing) { return state(packing).handle; }
void sync_mirror(const Packing &packing);
void mark_mirror_stale(const Packing &packing) {
    state(packing).mirrorStale = true;
    if (!state(packing).lazy) sync_mirror(packing);
}
bool set_lazy_mirror(const Packing &packing, bool lazy) {
    bool old = state(packing).lazy;
    state(packing).lazy = lazy;
    return old;
}

void boxToArray(const TriclinicBox &box, double out[9]) {
    const auto &dim = box.getDimensions();
    for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) out[3 * r + c] = dim(r, c);
}

}  // namespace rampack_b200

using namespace rampack_b200;

// host mirror <- device (everything, or nothing when it is current)
void rampack_b200::sync_mirror(const Packing &packing) {
    DeviceState &st = state(packing);
    if (!st.mirrorStale || packing.empty()) { st.mirrorStale = false; return; }
    auto &self = const_cast<Packing &>(packing);
    double box[9];
    check(st.handle, rp_download(st.handle, box, reinterpret_cast<double *>(self.shapes.data()), packing.size()));
    st.mirrorStale = false;
}

Packing::Packing(std::unique_ptr<BoundaryConditions> bc, std::size_t moveThreads, std::size_t scalingThreads)
        : bc{std::make_unique<DeviceState>(std::move(bc))}
{
    // the device sweeps need no OpenMP move threads; the temp slots are kept so that end() and the callers' thread
    // count checks (Simulation.cpp:64) behave as in the reference
    this->moveThreads = (moveThreads == 0 ? OMP_MAXTHREADS : moveThreads);
    this->scalingThreads = (scalingThreads == 0 ? OMP_MAXTHREADS : scalingThreads);
    this->shapes.resize(this->moveThreads);
    this->lastAlteredParticleIdx.resize(this->moveThreads, 0);
    this->lastMoveOverlapDeltas.resize(this->moveThreads, 0);
}

Packing::Packing(const TriclinicBox &box, std::vector<Shape> shapes, std::unique_ptr<BoundaryConditions> bc,
                 const Interaction &interaction, std::size_t moveThreads, std::size_t scalingThreads)
        : bc{std::make_unique<DeviceState>(std::move(bc))}
{
    this->moveThreads = (moveThreads == 0 ? OMP_MAXTHREADS : moveThreads);
    this->scalingThreads = (scalingThreads == 0 ? OMP_MAXTHREADS : scalingThreads);
    this->reset(std::move(shapes), box, interaction);
}

bool Packing::areShapesWithinBox(const std::vector<Shape> &shapes, const TriclinicBox &box) {
    for (const auto &shape : shapes)
        for (auto posCoord: box.absoluteToRelative(shape.getPosition()))
            if (posCoord < 0 || posCoord >= 1)
                return false;
    return true;
}

void Packing::reset(std::vector<Shape> newShapes, const TriclinicBox &newBox, const Interaction &newInteraction) {
    Expects(newBox.getVolume() != 0);
    Expects(newInteraction.getRangeRadius() > 0);
    Expects(!newShapes.empty());
    Expects(Packing::areShapesWithinBox(newShapes, newBox));

    DeviceState &st = state(*this);
    this->shapes = std::move(newShapes);
    this->box = newBox;
    this->interactionRange = newInteraction.getRangeRadius();
    std::size_t n = this->shapes.size();
    this->shapes.resize(n + this->moveThreads);
    this->lastAlteredParticleIdx.resize(this->moveThreads, 0);
    this->lastMoveOverlapDeltas.resize(this->moveThreads, 0);
    this->bc->setBox(this->box);

    rp_xc_program xc;
    rp_interaction_desc desc;
    if (!rampack_b200_describe(newInteraction, desc, xc))
        throw std::runtime_error("rampack_b200: this Interaction has no device implementation");
    st.xc = xc;
    st.desc = desc;
    if (st.desc.xc) st.desc.xc = &st.xc;
    st.hasDesc = true;
    check(st.handle, rp_set_interaction(st.handle, &st.desc));
    double boxArr[9];
    boxToArray(this->box, boxArr);
    check(st.handle, rp_upload(st.handle, boxArr, reinterpret_cast<const double *>(this->shapes.data()), n));
    st.uploaded = true;
    st.mirrorStale = false;
    this->numInteractionCentres = newInteraction.getInteractionCentres().size();
    this->rebuildNeighbourGrid();
}

void Packing::setupForInteraction(const Interaction &interaction) {
    DeviceState &st = state(*this);
    rp_xc_program xc;
    rp_interaction_desc desc;
    if (!rampack_b200_describe(interaction, desc, xc))
        throw std::runtime_error("rampack_b200: this Interaction has no device implementation");
    this->interactionRange = interaction.getRangeRadius();
    this->numInteractionCentres = interaction.getInteractionCentres().size();
    // the device keys its descriptor on the contents: Simulation calls this at the start of every run, mostly with
    // the interaction the packing already has
    desc.xc = nullptr;
    rp_interaction_desc current = st.desc;
    current.xc = nullptr;
    bool same = st.hasDesc && std::memcmp(&desc, &current, sizeof(desc)) == 0
                && std::memcmp(&xc, &st.xc, sizeof(xc)) == 0;
    if (!same) {
        st.xc = xc;
        st.desc = desc;
        if (desc.hard_kind == RP_HARD_XENOCOLLIDE) st.desc.xc = &st.xc;
        st.hasDesc = true;
        check(st.handle, rp_set_interaction(st.handle, &st.desc));
    }
    this->rebuildNeighbourGrid();
}

// The host keeps a NeighbourGrid object only for its GEOMETRY (getNeighbourGridCellDivisions is inline in the header
// and read by Simulation / CasinoMode); the cell lists live on the device (Packing.cpp:858-899 cell-size policy).
void Packing::rebuildNeighbourGrid() {
    if (this->empty()) return;
    double cellSize = this->interactionRange;
    double minCellSize = std::cbrt(this->getVolume() / static_cast<double>(this->size())) / 5;
    if (this->interactionRange < minCellSize)
        cellSize = minCellSize;
    auto heights = this->box.getHeights();
    this->neighbourGrid = std::nullopt;
    if (cellSize * 4 > *std::max_element(heights.begin(), heights.end())) return;
    if (cellSize * (1 + 1e-12) > *std::min_element(heights.begin(), heights.end())) return;
    this->neighbourGrid.emplace(this->box, cellSize, 1);
    this->neighbourGridRebuilds++;
}

const Shape &Packing::operator[](std::size_t i) const {
    Expects(i < this->size());
    sync_mirror(*this);
    return this->shapes[i];
}

const Shape &Packing::front() const {
    Expects(!this->empty());
    sync_mirror(*this);
    return this->shapes.front();
}

const Shape &Packing::back() const {
    Expects(!this->empty());
    sync_mirror(*this);
    return this->shapes[this->size() - 1];
}

namespace {
double explicitMove(Packing &packing, std::size_t particleIdx, int kind, const Vector<3> &translation,
                    const Matrix<3, 3> &rotation) {
    DeviceState &st = state(packing);
    rp_move m{};
    m.particle = particleIdx;
    m.kind = kind;
    for (int c = 0; c < 3; c++) m.translation[c] = translation[c];
    for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) m.rotation[3 * r + c] = rotation(r, c);
    double dE{};
    check(st.handle, rp_try_move(st.handle, &m, &dE));
    return dE;
}
}  // namespace

double Packing::tryTranslation(std::size_t particleIdx, Vector<3> translation, const Interaction &interaction,
                               std::optional<ActiveDomain> boundaries)
{
    Expects(particleIdx < this->size());
    Expects(interaction.getRangeRadius() <= this->interactionRange);
    this->lastAlteredParticleIdx[0] = particleIdx;
    if (boundaries.has_value()) {
        // ActiveDomain::isInside on the translated, wrapped mass centre (Packing.cpp:90-96)
        sync_mirror(*this);
        Shape temp = this->shapes[particleIdx];
        temp.translate(translation, *this->bc);
        if (!boundaries->isInside(temp.getPosition()))
            return std::numeric_limits<double>::infinity();
    }
    return explicitMove(*this, particleIdx, RP_MOVE_TRANSLATION, translation, Matrix<3, 3>::identity());
}

double Packing::tryRotation(std::size_t particleIdx, const Matrix<3, 3> &rotation, const Interaction &interaction) {
    Expects(particleIdx < this->size());
    Expects(interaction.getRangeRadius() <= this->interactionRange);
    this->lastAlteredParticleIdx[0] = particleIdx;
    return explicitMove(*this, particleIdx, RP_MOVE_ROTATION, Vector<3>{}, rotation);
}

double Packing::tryMove(std::size_t particleIdx, const Vector<3> &translation, const Matrix<3, 3> &rotation,
                        const Interaction &interaction, std::optional<ActiveDomain> boundaries)
{
    Expects(particleIdx < this->size());
    Expects(interaction.getRangeRadius() <= this->interactionRange);
    this->lastAlteredParticleIdx[0] = particleIdx;
    if (boundaries.has_value()) {
        sync_mirror(*this);
        Shape temp = this->shapes[particleIdx];
        temp.translate(translation, *this->bc);
        if (!boundaries->isInside(temp.getPosition()))
            return std::numeric_limits<double>::infinity();
    }
    return explicitMove(*this, particleIdx, RP_MOVE_ROTOTRANSLATION, translation, rotation);
}

void Packing::acceptMove() {
    DeviceState &st = state(*this);
    check(st.handle, rp_accept_move(st.handle));
    // one particle changed: refresh its mirror entry instead of declaring the whole mirror stale
    if (!st.mirrorStale) {
        std::size_t idx = this->lastAlteredParticleIdx[0];
        check(st.handle, rp_download_range(st.handle, idx, 1, reinterpret_cast<double *>(&this->shapes[idx])));
    }
}
void Packing::acceptTranslation() { this->acceptMove(); }
void Packing::acceptRotation() { this->acceptMove(); }

double Packing::tryScaling(const TriclinicBox &newBox, const Interaction &interaction) {
    Expects(interaction.getRangeRadius() <= this->interactionRange);
    DeviceState &st = state(*this);
    double boxArr[9], dE{};
    boxToArray(newBox, boxArr);
    check(st.handle, rp_try_scaling(st.handle, boxArr, &dE));
    this->lastBox = this->box;
    this->box = newBox;
    this->bc->setBox(this->box);
    mark_mirror_stale(*this);
    return dE;
}

double Packing::tryScaling(const std::array<double, 3> &scaleFactor, const Interaction &interaction) {
    Expects(std::all_of(scaleFactor.begin(), scaleFactor.end(), [](double d) { return d > 0; }));
    TriclinicBox newBox = this->box;
    newBox.scale(scaleFactor);
    return this->tryScaling(newBox, interaction);
}

void Packing::revertScaling() {
    DeviceState &st = state(*this);
    check(st.handle, rp_revert_scaling(st.handle));
    this->box = this->lastBox;
    this->bc->setBox(this->box);
    mark_mirror_stale(*this);
}

std::size_t Packing::countTotalOverlaps(const Interaction &interaction, bool earlyExit) const {
    if (!interaction.hasHardPart()) return 0;
    DeviceState &st = state(*this);
    uint64_t count{};
    check(st.handle, rp_count_total_overlaps(st.handle, earlyExit ? 1 : 0, &count));
    return count;
}

std::size_t Packing::countWallOverlaps(const Interaction &interaction, bool earlyExit) const {
    if (!this->hasAnyWalls || !interaction.hasWallPart()) return 0;
    DeviceState &st = state(*this);
    uint64_t count{};
    check(st.handle, rp_count_wall_overlaps(st.handle, earlyExit ? 1 : 0, &count));
    return count;
}

double Packing::getTotalEnergy(const Interaction &interaction) const {
    if (!interaction.hasSoftPart()) return 0;
    DeviceState &st = state(*this);
    double energy{};
    check(st.handle, rp_total_energy(st.handle, &energy));
    return energy;
}

double Packing::getParticleEnergyFluctuations(const Interaction &interaction) const {
    if (!interaction.hasSoftPart()) return 0;
    DeviceState &st = state(*this);
    double value{};
    check(st.handle, rp_particle_energy_fluctuations(st.handle, &value));
    return value;
}

void Packing::toggleOverlapCounting(bool countOverlaps, const Interaction &interaction) {
    DeviceState &st = state(*this);
    this->overlapCounting = countOverlaps;
    (void) interaction;      // the device counts with the interaction it was set up for (setupForInteraction)
    check(st.handle, rp_toggle_overlap_counting(st.handle, countOverlaps ? 1 : 0));
}

std::size_t Packing::getCachedNumberOfOverlaps() const {
    DeviceState &st = state(*this);
    uint64_t count{};
    rp_status status = rp_cached_overlaps(st.handle, &count);
    if (status == RP_ERR_STATE) throw std::runtime_error(rp_last_error(st.handle));     // Packing.cpp:1059
    check(st.handle, status);
    return count;
}

void Packing::toggleWall(std::size_t wallAxis, bool trueOrFalse) {
    Expects(wallAxis < 3);
    auto walls = this->hasWall;
    walls[wallAxis] = trueOrFalse;
    this->toggleWalls(walls);
}

void Packing::toggleWalls(std::array<bool, 3> axisWalls) {
    this->hasWall = axisWalls;
    this->hasAnyWalls = std::find(this->hasWall.begin(), this->hasWall.end(), true) != this->hasWall.end();
    DeviceState &st = state(*this);
    int32_t walls[3] = {axisWalls[0] ? 1 : 0, axisWalls[1] ? 1 : 0, axisWalls[2] ? 1 : 0};
    check(st.handle, rp_toggle_walls(st.handle, walls));
}

std::size_t Packing::renormalizeOrientations(const Interaction &interaction, bool allowOverlaps) {
    (void) interaction;
    DeviceState &st = state(*this);
    uint64_t rejected{};
    check(st.handle, rp_renormalize_orientations(st.handle, allowOverlaps ? 1 : 0, &rejected));
    mark_mirror_stale(*this);
    return rejected;
}

double Packing::getVolume() const { return std::abs(this->box.getVolume()); }

double Packing::getPackingFraction(double shapeVolume) const {
    Expects(shapeVolume >= 0);
    return this->getNumberDensity() * shapeVolume;
}

double Packing::getNumberDensity() const {
    return static_cast<double>(this->size()) / this->getVolume();
}

double Packing::getAverageNumberOfNeighbours() const {
    DeviceState &st = state(*this);
    double candidates{}, within{};
    check(st.handle, rp_neighbour_stats(st.handle, &candidates, &within));
    return candidates;
}

void Packing::store(std::ostream &out, const std::map<std::string, std::string> &auxInfo) const {
    sync_mirror(*this);
    RamsnapWriter writer;
    writer.write(out, *this, auxInfo);
}

std::map<std::string, std::string> Packing::restore(std::istream &in, const Interaction &interaction) {
    RamsnapReader reader;
    return reader.read(in, *this, interaction);
}

void Packing::resetCounters() {
    this->neighbourGridRebuilds = 0;
    this->neighbourGridResizes = 0;
    this->neighbourGridRebuildMicroseconds = 0;
}

std::ostream &operator<<(std::ostream &out, const Packing &packing) {
    sync_mirror(packing);
    out << "Packing {" << std::endl;
    out << "  box: {" << packing.box.getDimensions() << "}," << std::endl;
    out << "  particles (" << packing.size() << "): {" << std::endl;
    for (const auto &shape : packing)
        out << "    " << shape << std::endl;
    out << "  }" << std::endl;
    out << "}";
    return out;
}

std::size_t Packing::getShapesMemoryUsage() const { return get_vector_memory_usage(this->shapes); }

std::size_t Packing::getNeighbourGridMemoryUsage() const {
    return this->neighbourGrid.has_value() ? this->neighbourGrid->getMemoryUsage() : 0;
}

void Packing::resetNGRaceConditionSanitizer() { }

std::vector<Vector<3>> Packing::dumpNamedPoints(const ShapeGeometry &geometry, const std::string &pointName) const {
    sync_mirror(*this);
    std::vector<Vector<3>> namedPoints;
    namedPoints.reserve(this->size());
    Vector<3> alignedNamedPoint = geometry.getNamedPoint(pointName);
    for (const auto &shape : *this)
        namedPoints.push_back(shape.getPosition() + shape.getOrientation() * alignedNamedPoint);
    return namedPoints;
}
