// flat_scene.h -- host-side flat scene (the output of the scene compiler, the input of the uploader).
#pragma once
#include "../../../include/b200pt.h"
#include "json.h"

#include <memory>
#include <string>
#include <vector>

namespace b200pt {

struct Aabb {
    float lo[3], hi[3];
};

struct ImageData { // image_texture (image.h:7-70)
    int32_t width = 0, height = 0;
    std::vector<float> rgba; // 4 * width * height
};

struct BuiltBvh;

struct FlatScene {
    std::vector<b200pt_instance> instances;
    std::vector<b200pt_material> materials;
    std::vector<b200pt_texture> textures;
    std::vector<int32_t> lights;
    int32_t background = 0;
    b200pt_camera camera{};
    std::vector<float> perlin_ranvec;  // 256*3 (always filled: cheap)
    std::vector<int32_t> perlin_perm;  // 3*256
    std::vector<ImageData> images;     // B200PT_TEX_IMAGE textures refer to these by index (texture.odd)
    mutable std::vector<b200pt_image> image_views; // desc() scratch: views into `images`
    std::vector<std::string> warnings; // parser warnings the reference prints to stdout
    // compile_scene builds the tree of a large scene on a second thread while it replays the reference's instance
    // order; pack_scene takes it from here (null: pack_scene builds it)
    std::shared_ptr<const BuiltBvh> prebuilt_bvh;

    b200pt_scene_desc desc() const;
    static FlatScene from_desc(const b200pt_scene_desc &d);
};

// build_scene (scene_parser.h:241-595) + setup_camera (main.cpp:86-104).  Throws std::runtime_error.
// asset_dir: where relative "path" entries of png textures are looked up after the working directory (the reference
// opens them relative to the working directory only, scene_parser.h:39-55); may be empty.
FlatScene compile_scene(const Json &scene, int film_width, int film_height, const std::string &asset_dir = "");

// 8/16-bit PNG (grey, grey+alpha, RGB, palette, RGBA; non-interlaced) -> RGBA8 like lodepng::decode's default.
// Throws std::runtime_error.
void decode_png_rgba8(const std::string &path, std::vector<unsigned char> &rgba, unsigned &width, unsigned &height);

// camera::camera (camera.h:9-36) in the reference's float arithmetic.
b200pt_camera make_camera(const float lookfrom[3], const float lookat[3], float vfov_deg, float aspect, float aperture,
                          float focus_dist);

// transform3(scale, rotate, translate) (transform3.h:17-25): T * Rx * Ry * Rz * S, angles in units of pi, composed
// through unit quaternions exactly as Eigen 3.2.10 does for AngleAxisf products, in float.
void make_transform(const float scale[3], const float rotate[3], const float translate[3], float out[12]);

// Inverse of a 3x4 affine exactly as Eigen 3.2.10 computes Affine3f::inverse() in float (the reference calls it on
// every hit, transform3.h:51-54).  Returns false when singular.
bool invert_affine(const float m[12], float out[12]);

// instance::instance bbox (primitive.h:266-297): the primitive's own box, 8 corners through the transform.
Aabb instance_bounds(const b200pt_instance &inst);

// The reference draws its Perlin tables (texture.h:76-110,180-183) and its BVH split axes (bvh.h:135) from
// one default-seeded std::mt19937 (random.h:9-15).  This replays that stream: fills the tables, then walks the
// same recursive median-split build over the instance list to find which instances end up alone in an n==1
// leaf (bvh.h:150-153), where the reference intersects them twice, and REORDERS `instances` into the left-to-right
// leaf order of that tree (lights[] is re-indexed), so that the reference's tie rule for coincident surfaces
// (bvh.h:57-60: the right child wins) becomes "the higher instance index wins".
void replay_reference_rng(FlatScene &scene);

} // namespace b200pt
