// host/describe_interaction.h — see describe_interaction.cpp
#pragma once

#include "rampack_b200.h"

class Interaction;

// Fills `desc` (and, for XenoCollide shapes, `xc`, which desc.xc then points to) from a reference Interaction object.
// Returns false when the interaction has no device counterpart.
bool rampack_b200_describe(const Interaction &interaction, rp_interaction_desc &desc, rp_xc_program &xc);
