// host/describe_interaction.cpp — reference-side glue: an `Interaction` of the reference as the POD the device wants.
//
// The reference hides shape parameters behind a vtable and private members (e.g. SpherocylinderTraits has no getters
// for length / radius, src/core/shapes/SpherocylinderTraits.h:27-28).  A maintainer would add one virtual
// `describe()` to Interaction; this translation unit does the same job WITHOUT touching the
// reference's sources: it is compiled with -fno-access-control against the unmodified headers and reads the
// parameters with dynamic_cast.  It contains no arithmetic of its own.
//
// Covered: SphereTraits::HardInteraction, SpherocylinderTraits, PolysphereTraits::HardInteraction (kmer, banana,
// lollipop, wedge ... all derive their interaction from it), PolyspherocylinderTraits, the CentralInteraction
// family on centres (LennardJones / RepulsiveLennardJones / SquareInverseCore), CompoundInteraction (hard + soft),
// XenoCollideTraits<SmoothWedgeTraits | PolyhedralWedgeTraits | GenericXenoCollideTraits> with every primitive and
// operation of geometry/xenocollide/.
#include <cstring>
#include <map>
#include <stdexcept>
#include <string>

#include "core/Interaction.h"
#include "core/shapes/SphereTraits.h"
#include "core/shapes/SpherocylinderTraits.h"
#include "core/shapes/PolysphereTraits.h"
#include "core/shapes/PolyspherocylinderTraits.h"
#include "core/shapes/SmoothWedgeTraits.h"
#include "core/shapes/PolyhedralWedgeTraits.h"
#include "core/shapes/GenericXenoCollideTraits.h"
#include "core/interactions/CentralInteraction.h"
#include "core/interactions/LennardJonesInteraction.h"
#include "core/interactions/RepulsiveLennardJonesInteraction.h"
#include "core/interactions/SquareInverseCoreInteraction.h"
#include "core/interactions/CompoundInteraction.h"
#include "geometry/xenocollide/XCPrimitives.h"
#include "geometry/xenocollide/XCOperations.h"

#include "describe_interaction.h"

namespace {

void setCentres(rp_interaction_desc &d, const std::vector<Vector<3>> &centres) {
    if (centres.size() > RP_MAX_CENTRES)
        throw std::runtime_error("rampack_b200: more than " + std::to_string(RP_MAX_CENTRES) + " interaction centres");
    d.n_centres = static_cast<int32_t>(centres.size());
    for (std::size_t i{}; i < centres.size(); i++)
        for (int c = 0; c < 3; c++) d.centres[i][c] = centres[i][c];
}

bool describeSoft(const Interaction &I, rp_interaction_desc &d) {
    if (auto lj = dynamic_cast<const LennardJonesInteraction *>(&I)) {
        d.soft_kind = RP_SOFT_LJ; d.epsilon = lj->epsilon; d.sigma = lj->sigma;
    } else if (auto wca = dynamic_cast<const RepulsiveLennardJonesInteraction *>(&I)) {
        d.soft_kind = RP_SOFT_WCA; d.epsilon = wca->epsilon; d.sigma = wca->sigma;
    } else if (auto sic = dynamic_cast<const SquareInverseCoreInteraction *>(&I)) {
        d.soft_kind = RP_SOFT_SIC; d.epsilon = sic->epsilon; d.sigma = sic->sigma;
    } else {
        return false;
    }
    setCentres(d, I.getInteractionCentres());
    return true;
}

// ---- XenoCollide geometry tree -> flat program (children before parents) ----
struct XcWalk {
    rp_xc_program &P;
    std::map<const AbstractXCGeometry *, int> seen;

    int leaf(int kind, std::initializer_list<double> params) {
        if (P.n_nodes >= RP_XC_MAX_NODES) throw std::runtime_error("rampack_b200: XenoCollide geometry has too many nodes");
        rp_xc_node &n = P.nodes[P.n_nodes];
        n.kind = kind;
        int i = 0;
        for (double v : params) n.params[i++] = v;
        return P.n_nodes++;
    }

    template <typename Entries>
    int operation(int kind, const Entries &entries) {
        std::vector<int> children;
        for (const auto &e : entries) children.push_back(this->walk(*e.geometry));
        if (P.n_nodes >= RP_XC_MAX_NODES || P.n_entries + static_cast<int>(children.size()) > RP_XC_MAX_ENTRIES)
            throw std::runtime_error("rampack_b200: XenoCollide geometry has too many nodes / entries");
        rp_xc_node &n = P.nodes[P.n_nodes];
        n.kind = kind;
        n.n_entries = static_cast<int32_t>(children.size());
        n.first_entry = P.n_entries;
        std::size_t k{};
        for (const auto &e : entries) {
            rp_xc_entry &out = P.entries[P.n_entries++];
            out.node = children[k++];
            for (int c = 0; c < 3; c++) out.pos[c] = e.pos[c];
            for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) out.rot[3 * r + c] = e.rot(r, c);
        }
        return P.n_nodes++;
    }

    int walk(const AbstractXCGeometry &g) {
        auto it = seen.find(&g);
        if (it != seen.end()) return it->second;
        int idx;
        if (auto p = dynamic_cast<const XCPoint *>(&g)) idx = leaf(RP_XC_POINT, {p->pos[0], p->pos[1], p->pos[2]});
        else if (auto s = dynamic_cast<const XCSegment *>(&g)) idx = leaf(RP_XC_SEGMENT, {s->halfLength});
        else if (auto r = dynamic_cast<const XCRectangle *>(&g)) idx = leaf(RP_XC_RECTANGLE, {r->halfSides[0], r->halfSides[1]});
        else if (auto c = dynamic_cast<const XCCuboid *>(&g)) idx = leaf(RP_XC_CUBOID, {c->halfSides[0], c->halfSides[1], c->halfSides[2]});
        else if (auto dk = dynamic_cast<const XCDisk *>(&g)) idx = leaf(RP_XC_DISK, {dk->radius});
        else if (auto sp = dynamic_cast<const XCSphere *>(&g)) idx = leaf(RP_XC_SPHERE, {sp->radius});
        else if (auto e = dynamic_cast<const XCEllipse *>(&g)) idx = leaf(RP_XC_ELLIPSE, {e->semiAxes[0], e->semiAxes[1]});
        else if (auto el = dynamic_cast<const XCEllipsoid *>(&g)) idx = leaf(RP_XC_ELLIPSOID, {el->semiAxes[0], el->semiAxes[1], el->semiAxes[2]});
        else if (auto f = dynamic_cast<const XCFootball *>(&g)) idx = leaf(RP_XC_FOOTBALL, {f->halfLength, f->radius});
        else if (auto sa = dynamic_cast<const XCSaucer *>(&g)) idx = leaf(RP_XC_SAUCER, {sa->radius, sa->halfThickness});
        else if (auto sum = dynamic_cast<const XCSum *>(&g)) idx = operation(RP_XC_SUM, sum->entries);
        else if (auto mx = dynamic_cast<const XCMax *>(&g)) idx = operation(RP_XC_MAX, mx->entries);
        else if (auto df = dynamic_cast<const XCDiff *>(&g)) {
            struct E { std::shared_ptr<AbstractXCGeometry> geometry; Vector<3> pos; Matrix<3, 3> rot; };
            std::vector<E> entries = {{df->geom1, df->pos1, df->rot1}, {df->geom2, df->pos2, df->rot2}};
            idx = operation(RP_XC_DIFF, entries);
        } else {
            throw std::runtime_error("rampack_b200: unknown AbstractXCGeometry subclass");
        }
        seen[&g] = idx;
        return idx;
    }
};

bool describeXenoCollide(const Interaction &I, rp_interaction_desc &d, rp_xc_program &P) {
    std::memset(&P, 0, sizeof(P));
    if (auto sw = dynamic_cast<const SmoothWedgeTraits *>(&I)) {
        setCentres(d, sw->interactionCentres);
        if (sw->shapeModel.size() > RP_MAX_CENTRES) return false;
        for (std::size_t i{}; i < sw->shapeModel.size(); i++) {
            const auto &g = sw->shapeModel[i];
            rp_xc_node &n = P.nodes[P.n_nodes];
            n.kind = RP_XC_SMOOTH_WEDGE;
            n.params[0] = g.R; n.params[1] = g.r; n.params[2] = g.l;
            P.root[i] = P.n_nodes++;
        }
    } else if (auto pw = dynamic_cast<const PolyhedralWedgeTraits *>(&I)) {
        setCentres(d, pw->interactionCentres);
        if (pw->shapeModels.size() > RP_MAX_CENTRES) return false;
        for (std::size_t i{}; i < pw->shapeModels.size(); i++) {
            // CollideGeometry(axBottom, ayBottom, axTop, ayTop, length) keeps the two half-size vertices
            // (PolyhedralWedgeTraits.cpp): vertexUp = (axTop / 2, ayTop / 2, length / 2), vertexDown = bottom
            const auto &g = pw->shapeModels[i];
            rp_xc_node &n = P.nodes[P.n_nodes];
            n.kind = RP_XC_POLYHEDRAL_WEDGE;
            n.params[0] = 2 * g.vertexDown[0]; n.params[1] = 2 * g.vertexDown[1];
            n.params[2] = 2 * g.vertexUp[0]; n.params[3] = 2 * g.vertexUp[1];
            n.params[4] = 2 * g.vertexUp[2];
            P.root[i] = P.n_nodes++;
        }
    } else if (auto gen = dynamic_cast<const GenericXenoCollideTraits *>(&I)) {
        XcWalk walk{P, {}};
        P.root[0] = walk.walk(*gen->geometry);
        d.n_centres = 0;
    } else {
        return false;
    }
    d.hard_kind = RP_HARD_XENOCOLLIDE;
    d.xc = &P;
    return true;
}

bool describeHard(const Interaction &I, rp_interaction_desc &d, rp_xc_program &xc) {
    if (auto s = dynamic_cast<const SphereTraits::HardInteraction *>(&I)) {
        d.hard_kind = RP_HARD_SPHERE; d.radius = s->radius;
        return true;
    }
    if (auto sc = dynamic_cast<const SpherocylinderTraits *>(&I)) {
        d.hard_kind = RP_HARD_SPHEROCYLINDER; d.radius = sc->radius; d.length = sc->length;
        return true;
    }
    if (auto ps = dynamic_cast<const PolysphereTraits::HardInteraction *>(&I)) {
        d.hard_kind = RP_HARD_POLYSPHERE;
        setCentres(d, I.getInteractionCentres());
        for (std::size_t i{}; i < ps->sphereData.size(); i++) d.centre_radius[i] = ps->sphereData[i].radius;
        return true;
    }
    if (auto psc = dynamic_cast<const PolyspherocylinderTraits *>(&I)) {
        d.hard_kind = RP_HARD_POLYSPHEROCYLINDER;
        setCentres(d, I.getInteractionCentres());
        const auto &data = psc->getSpherocylinderData();
        for (std::size_t i{}; i < data.size(); i++) {
            d.centre_radius[i] = data[i].radius;
            for (int c = 0; c < 3; c++) d.centre_half_axis[i][c] = data[i].halfAxis[c];
        }
        return true;
    }
    return describeXenoCollide(I, d, xc);
}

}  // namespace

bool rampack_b200_describe(const Interaction &interaction, rp_interaction_desc &d, rp_xc_program &xc) {
    std::memset(&d, 0, sizeof(d));
    if (auto compound = dynamic_cast<const CompoundInteraction *>(&interaction)) {
        // CompoundInteraction.cpp:5-66: OR of the hard parts, sum of the soft parts; the device handles one hard and
        // one soft constituent (what overlap relaxation builds, frontend CasinoMode.cpp:288-289)
        const Interaction *parts[2] = {&compound->interaction1, &compound->interaction2};
        bool hard = false, soft = false;
        for (const Interaction *part : parts) {
            rp_interaction_desc one;
            std::memset(&one, 0, sizeof(one));
            if (part->hasHardPart() && !part->hasSoftPart() && !hard && describeHard(*part, one, xc)) {
                hard = true;
                d.hard_kind = one.hard_kind; d.radius = one.radius; d.length = one.length; d.xc = one.xc;
                std::memcpy(d.centre_radius, one.centre_radius, sizeof(one.centre_radius));
                std::memcpy(d.centre_half_axis, one.centre_half_axis, sizeof(one.centre_half_axis));
            } else if (part->hasSoftPart() && !part->hasHardPart() && !soft && describeSoft(*part, one)) {
                soft = true;
                d.soft_kind = one.soft_kind; d.epsilon = one.epsilon; d.sigma = one.sigma;
            } else {
                return false;
            }
        }
        setCentres(d, interaction.getInteractionCentres());
        return hard && soft;
    }
    if (interaction.hasHardPart() && !interaction.hasSoftPart()) return describeHard(interaction, d, xc);
    if (interaction.hasSoftPart() && !interaction.hasHardPart()) return describeSoft(interaction, d);
    return false;
}
