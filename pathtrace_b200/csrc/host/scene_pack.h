// scene_pack.h -- FlatScene -> the arrays the kernels read (PtInstance / PtFwd / PtNode4), still in host memory.
#pragma once
#include "../device/pt_scene.h"
#include "bvh_build.h"
#include "flat_scene.h"

#include <vector>

namespace b200pt {

struct PackedScene {
    std::vector<pt::PtInstance> inst;
    std::vector<pt::PtFwd> fwd;
    BuiltBvh bvh;
    std::vector<float> perlin_ranvec;  // 768
    std::vector<int32_t> perlin_perm;  // 768
    std::vector<pt::PtBox> small; // filled when the scene has <= PT_SMALL_MAX instances (B200PT_SMALL_MAX=0: never)
    std::vector<pt::PtImage> images;     // image textures: dimensions + offset into texels
    std::vector<pt::PtTexel> texels;
    std::vector<int32_t> emitters;     // instances whose hit material is a diffuse_light, ascending
    std::vector<pt::PtBox> light_boxes; // world bounds per light (flat.lights order); empty: the horizon test is off
    int n_media = 0;
    float sort_bounds[6] = {0, 0, 0, 1, 1, 1}; // bounds of the non-medium instances (coherence sort cells)
    // a view whose pointers refer to THIS object's host vectors and to `flat`'s material/texture/light arrays
    pt::PtSceneView host_view(const FlatScene &flat) const;
};

PackedScene pack_scene(const FlatScene &flat);

} // namespace b200pt
