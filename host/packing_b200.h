// host/packing_b200.h — what the Simulation hooks (Simulation_b200.cpp) need from the Packing facade
#pragma once

#include "rampack_b200.h"

class Packing;

namespace rampack_b200 {
    rp_handle *handle_of(const Packing &packing);
    // brings Packing::shapes up to date with the device (no-op when it is); observers read the mirror through the
    // inline begin() / end() of the unmodified header, so the cycle loop calls this before anybody may look
    void sync_mirror(const Packing &packing);
    void mark_mirror_stale(const Packing &packing);
    // lazy = true: mutating calls leave the mirror stale until sync_mirror; returns the previous mode
    bool set_lazy_mirror(const Packing &packing, bool lazy);
    void check(rp_handle *h, rp_status st);
}
