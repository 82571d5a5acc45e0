#pragma once
#include "../device/pt_scene.h"
#include "flat_scene.h"

#include <vector>

namespace b200pt {

struct BuiltBvh {
    std::vector<pt::PtNode> nodes; // node 0 is the root
    std::vector<pt::PtNode4> nodes4; // the same tree, up to 4 children per node (what the kernels walk)
    int max_depth4 = 0;
    int max_stack4 = 0; // pending entries a traversal of nodes4 can need at most
    int max_depth = 0;
    int n_leaves = 0; // instances the tree was built over
    float bounds[6] = {0, 0, 0, 0, 0, 0};
};

BuiltBvh build_bvh(const std::vector<Aabb> &boxes);

} // namespace b200pt
