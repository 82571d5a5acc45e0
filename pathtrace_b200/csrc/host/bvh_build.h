#pragma once
#include "../device/pt_scene.h"
#include "flat_scene.h"

#include <vector>

namespace b200pt {

struct BuiltBvh {
    std::vector<pt::PtNode> nodes; // node 0 is the root
    int max_depth = 0;
    float bounds[6] = {0, 0, 0, 0, 0, 0};
};

BuiltBvh build_bvh(const std::vector<Aabb> &boxes);

} // namespace b200pt
