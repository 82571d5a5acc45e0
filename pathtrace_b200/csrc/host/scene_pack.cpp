#include "scene_pack.h"

#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <stdexcept>
#include <string>

namespace b200pt {

static pt::V3 to_v3(const float *p) { return pt::v3(p[0], p[1], p[2]); }

PackedScene pack_scene(const FlatScene &f)
{
    PackedScene out;
    const int n = (int)f.instances.size();
    if (n <= 0)
        throw std::runtime_error("scene has no instances");
    if (n >= (1 << 28))
        throw std::runtime_error("too many instances (28-bit instance index in the wavefront hit record)");
    if (f.materials.size() > 32767)
        throw std::runtime_error("too many materials (15-bit material index)");
    out.inst.resize(n);
    out.fwd.resize(n);
    std::vector<Aabb> boxes(n);
    int media = 0;
    for (int i = 0; i < n; i++) {
        const b200pt_instance &src = f.instances[i];
        pt::PtInstance &dst = out.inst[i];
        // transform3::inverse() is re-derived by the reference on every call (transform3.h:51-54); here once
        if (!invert_affine(src.fwd, dst.inv))
            throw std::runtime_error("instance " + std::to_string(i) + " has a singular transform");
        memcpy(out.fwd[i].m, src.fwd, sizeof(src.fwd));
        uint32_t meta = (uint32_t)(src.prim_kind & 3) | ((uint32_t)(src.plane & 3) << 2);
        if (src.flags & B200PT_INST_RECT_FLIP)
            meta |= 1u << 4;
        if (src.flags & B200PT_INST_MEDIUM) {
            if (media >= 256)
                throw std::runtime_error("more than 256 media in one scene");
            meta |= 1u << 5;
            meta |= (uint32_t)media << 8;
            media++;
        }
        if (src.flags & B200PT_INST_BVH_DUP)
            meta |= 1u << 6;
        if (src.light_index >= 0)
            meta |= 1u << 7;
        meta |= (uint32_t)src.material << 16;
        if (src.material >= 0 && src.material < (int)f.materials.size() &&
            f.materials[src.material].kind == B200PT_MAT_DIFFUSE_LIGHT) {
            meta |= 1u << 31;
            out.emitters.push_back(i);
        }
        dst.meta = meta;
        memcpy(dst.p, src.params, sizeof(dst.p));
        dst.density = src.density;
        boxes[i] = instance_bounds(src);
    }
    out.n_media = media;
    {
        // cells of the coherence sort: the bounds of everything that is not a medium (all instances if nothing else)
        float lo[3] = {3e38f, 3e38f, 3e38f}, hi[3] = {-3e38f, -3e38f, -3e38f};
        for (int pass = 0; pass < 2 && !(lo[0] <= hi[0]); pass++)
            for (int i = 0; i < n; i++) {
                if (pass == 0 && (f.instances[i].flags & B200PT_INST_MEDIUM))
                    continue;
                for (int k = 0; k < 3; k++) {
                    if (boxes[i].lo[k] > -1e30f && boxes[i].lo[k] < lo[k])
                        lo[k] = boxes[i].lo[k];
                    if (boxes[i].hi[k] < 1e30f && boxes[i].hi[k] > hi[k])
                        hi[k] = boxes[i].hi[k];
                }
            }
        for (int k = 0; k < 3; k++) {
            out.sort_bounds[k] = lo[k] <= hi[k] ? lo[k] : 0.f;
            out.sort_bounds[3 + k] = lo[k] <= hi[k] ? hi[k] : 1.f;
        }
    }
    // World bounds of the lights, for the horizon test of pt_shade.h lights_below_horizon: only lights whose samples
    // point AT them (rect::random, sphere::random); a box or medium in the light list samples the fixed direction
    // (1, 0, 0) (hittable.h:28) and switches the test off.
    if (!f.lights.empty() && f.lights.size() <= PT_MAX_LIGHT_BOXES) {
        bool ok = true;
        for (int32_t li : f.lights) {
            const b200pt_instance &src = f.instances[li];
            if ((src.flags & B200PT_INST_MEDIUM) || (src.prim_kind != B200PT_PRIM_RECT && src.prim_kind != B200PT_PRIM_SPHERE))
                ok = false;
        }
        if (ok)
            for (int32_t li : f.lights) {
                pt::PtBox b{};
                for (int k = 0; k < 3; k++) {
                    b.lo[k] = boxes[li].lo[k];
                    b.hi[k] = boxes[li].hi[k];
                }
                out.light_boxes.push_back(b);
            }
    }
    for (const ImageData &im : f.images) {
        pt::PtImage pi{};
        pi.width = im.width;
        pi.height = im.height;
        pi.offset = (uint32_t)out.texels.size();
        out.images.push_back(pi);
        const size_t ntex = (size_t)im.width * im.height;
        for (size_t k = 0; k < ntex; k++)
            out.texels.push_back(pt::PtTexel{im.rgba[4 * k], im.rgba[4 * k + 1], im.rgba[4 * k + 2], im.rgba[4 * k + 3]});
    }
    if (f.prebuilt_bvh && f.prebuilt_bvh->n_leaves == n)
        out.bvh = *f.prebuilt_bvh;
    else
        out.bvh = build_bvh(boxes);
    if (out.bvh.max_stack4 > PT_STACK_DEPTH)
        throw std::runtime_error("BVH too deep for the traversal stack (" + std::to_string(out.bvh.max_stack4) + " > " +
                                 std::to_string(PT_STACK_DEPTH) + " pending nodes)");
    // B200PT_SMALL_MAX (tests, A/B): 0 forces the BVH path on small scenes too
    int small_max = PT_SMALL_MAX;
    if (const char *e = std::getenv("B200PT_SMALL_MAX"))
        small_max = std::min(PT_SMALL_MAX, std::max(0, atoi(e)));
    if (n <= small_max) {
        out.small.resize(n);
        for (int i = 0; i < n; i++) {
            pt::PtBox &e = out.small[i];
            for (int k = 0; k < 3; k++) {
                e.lo[k] = boxes[i].lo[k];
                e.hi[k] = boxes[i].hi[k];
            }
            e.pad0 = e.pad1 = 0.f;
        }
    }
    out.perlin_ranvec = f.perlin_ranvec;
    out.perlin_perm = f.perlin_perm;
    if (out.perlin_ranvec.size() != 768)
        out.perlin_ranvec.assign(768, 0.f);
    if (out.perlin_perm.size() != 768)
        out.perlin_perm.assign(768, 0);
    return out;
}

pt::PtSceneView PackedScene::host_view(const FlatScene &f) const
{
    pt::PtSceneView v{};
    v.nodes = bvh.nodes4.data();
    v.inst = inst.data();
    v.fwd = fwd.data();
    v.mats = f.materials.data();
    v.tex = f.textures.data();
    v.lights = f.lights.data();
    v.images = images.data();
    v.texels = texels.data();
    v.n_images = (int)images.size();
    v.emitters = emitters.data();
    v.light_boxes = light_boxes.data();
    v.n_light_boxes = (int)light_boxes.size();
    v.small = small.data();
    v.n_small = (int)small.size();
    v.n_emitters = (int)emitters.size();
    v.perlin_ranvec = perlin_ranvec.data();
    v.perlin_perm = perlin_perm.data();
    v.n_nodes = (int)bvh.nodes4.size();
    v.n_inst = (int)inst.size();
    v.n_lights = (int)f.lights.size();
    v.background = f.background;
    v.n_media = n_media;
    v.emit_needs_p = 0;
    v.emit_needs_n = 0;
    for (const b200pt_material &m : f.materials)
        if (m.kind == B200PT_MAT_DIFFUSE_LIGHT && !m.two_sided)
            v.emit_needs_n = 1;
    for (const b200pt_material &m : f.materials)
        if (m.kind == B200PT_MAT_DIFFUSE_LIGHT &&
            !(m.texture >= 0 && m.texture < (int)f.textures.size() && f.textures[m.texture].kind == B200PT_TEX_CONSTANT))
            v.emit_needs_p = 1;
    for (int k = 0; k < 3; k++) {
        const float ext = sort_bounds[3 + k] - sort_bounds[k];
        v.sort_lo[k] = sort_bounds[k];
        v.sort_scale[k] = ext > 0.f ? 1.0f / ext : 0.f; // position -> [0, 1) of the sort bounds
    }
    v.cam.origin = to_v3(f.camera.origin);
    v.cam.lower_left = to_v3(f.camera.lower_left);
    v.cam.horizontal = to_v3(f.camera.horizontal);
    v.cam.vertical = to_v3(f.camera.vertical);
    v.cam.u = to_v3(f.camera.u);
    v.cam.v = to_v3(f.camera.v);
    v.cam.lens_radius = f.camera.lens_radius;
    return v;
}

} // namespace b200pt
