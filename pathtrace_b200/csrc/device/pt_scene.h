// pt_scene.h -- the device-resident scene: plain structs, 16-byte aligned, uploaded once.
//
//   PtNode4    128 B  one node of the 4-wide BVH, child-major: {lo.x[4] | lo.y[4] | lo.z[4] | hi.x[4] | hi.y[4] | hi.z[4] |
//                     ref[4]}, fetched as 7 x 128-bit loads; 32 bytes of it per child a ray tests (slab test).
//   PtNode      64 B  host only: the binary SAH tree the builder makes before collapsing it.
//   PtInstance  80 B  world->local 3x4 (48 B) + tagged-union primitive (32 B): meta word, 6 params, density.
//   PtBox       32 B  small scenes only: the leaf box of every instance, walked linearly instead of the tree.
//   PtFwd       48 B  local->world 3x4, fetched only for the closest hit and for light sampling.
//   b200pt_material / b200pt_texture (32 B each) are used as they come from the C ABI.
#pragma once
#include "../../../include/b200pt.h"
#include "pt_math.h"

namespace pt {

struct alignas(16) PtNode {
    float lo0[3];
    int32_t ref0; // >= 0: inner node index; < 0: ~ref = instance index
    float hi0[3];
    int32_t pad0;
    float lo1[3];
    int32_t ref1;
    float hi1[3];
    int32_t pad1;
};

// The tree the kernels walk: the binary tree collapsed to 4 children per node (bvh_build.cpp collapse_node), stored child-
// major so that one 128-bit load brings the same bound of all four.  Half as many dependent node fetches per ray as
// the binary tree, and rays of a warp agree for longer on what to do next.
#define PT_NO_CHILD 0x7fffffff
#define PT_STACK_DEPTH 48 // traversal stack entries per ray; scene_pack.cpp refuses a tree that could need more
struct alignas(16) PtNode4 {
    float lox[4], loy[4], loz[4], hix[4], hiy[4], hiz[4];
    int32_t ref[4]; // >= 0: inner node index; < 0: ~ref = instance index; PT_NO_CHILD: empty slot
    int32_t pad[4];
};

// meta: bits 0-1 prim kind | 2-3 rect plane | 4 rect flip | 5 medium | 6 bvh-dup | 7 light | 8-15 medium slot |
//       16-30 material index | 31 emitter (the material a hit reports is a diffuse_light)
struct alignas(16) PtInstance {
    float inv[12];
    uint32_t meta;
    float p[6];
    float density;
};

// Small scenes (<= PT_SMALL_MAX instances; every Cornell-box config) skip the tree: all instances are walked in
// index order by the 32 lanes together (pt_intersect.h run_small), each behind the slab test of its leaf box
// (PtBox, 32 B).  The kernels keep boxes, instances and forward transforms of such a scene in shared memory.
#define PT_SMALL_MAX 16
#define PT_MAX_LIGHT_BOXES 8
struct alignas(16) PtBox {
    float lo[3];
    float pad0;
    float hi[3];
    float pad1;
};

struct alignas(16) PtFwd {
    float m[12];
};

PT_HD int meta_kind(uint32_t m) { return (int)(m & 3u); }
PT_HD int meta_plane(uint32_t m) { return (int)((m >> 2) & 3u); }
PT_HD bool meta_flip(uint32_t m) { return (m >> 4) & 1u; }
PT_HD bool meta_medium(uint32_t m) { return (m >> 5) & 1u; }
PT_HD bool meta_dup(uint32_t m) { return (m >> 6) & 1u; }
PT_HD bool meta_light(uint32_t m) { return (m >> 7) & 1u; }
PT_HD uint32_t meta_medium_slot(uint32_t m) { return (m >> 8) & 0xffu; }
PT_HD int meta_material(uint32_t m) { return (int)((m >> 16) & 0x7fffu); }
PT_HD bool meta_emitter(uint32_t m) { return (m >> 31) & 1u; }

// image_texture (image.h:7-51): texels of all images in one float4 array (rgb + alpha), row-major, first row of the
// file first
struct alignas(16) PtImage {
    int32_t width, height;
    uint32_t offset; // first texel in PtSceneView::texels
    int32_t pad;
};

struct alignas(16) PtTexel {
    float r, g, b, a;
};

#ifndef PT_SORT_CELL_BITS
#define PT_SORT_CELL_BITS 6 /* Morton bits per axis of the coherence-sort key (kernels.cu) */
#endif

struct PtCamera {
    V3 origin, lower_left, horizontal, vertical, u, v;
    float lens_radius;
};

struct PtSceneView {
    const PtNode4 *nodes;
    const PtInstance *inst;
    const PtFwd *fwd;
    const b200pt_material *mats;
    const b200pt_texture *tex;
    const int32_t *lights;
    const PtBox *small;         // n_small leaf boxes, index = instance index; n_small == 0: use the BVH
    const int32_t *emitters;    // instances whose hit material emits (shadow rays test these first, pt_intersect.h)
    const PtBox *light_boxes;   // world bounds of lights[0..n_light_boxes), see lights_below_horizon (pt_shade.h)
    const PtImage *images;
    const PtTexel *texels;
    const float *perlin_ranvec; // 256 x 3
    const int32_t *perlin_perm; // 3 x 256
    int32_t n_nodes, n_inst, n_lights, n_emitters, n_small, n_images, background;
    int32_t emit_needs_n; // some emitter is one-sided: shadow hits must report the normal (material.h:215-222)
    int32_t emit_needs_p; // some emitter has a non-constant texture: shadow hits must report p (else only the normal)
    int32_t n_media; // instances wrapped in a constant_medium; 0 selects the kernels compiled without that branch
    int32_t n_light_boxes; // n_lights when every light is a rect or a sphere (and there are <= 8), else 0
    // coherence sort keys (kernels.cu sort_key_*): cell = (p - sort_lo) * sort_scale * 2^bits, clamped; the bounds are those of
    // the instances that are not media (a scene-wide mist would make every cell cover the whole scene)
    float sort_lo[3], sort_scale[3];
    PtCamera cam;
};

} // namespace pt
