// pt_intersect.h -- closest-hit queries: primitive tests, the instance wrapper, the constant medium and the
// ordered stack traversal of the flat BVH.  Formula order follows the reference so that t / normal agree to
// fp32 rounding:
//   sphere::hit           primitive.h:64-95        rect::hit      primitive.h:186-225
//   box (6 rects, list)   primitive.h:233-246, hittable_list.h:21-38
//   constant_medium::hit  volume.h:29-93           instance::hit  primitive.h:298-312
//   aabb::hit             aabb.h:34-53             bvh_node::hit  bvh.h:31-69 (same closest hit, fewer visits)
#pragma once
#include "pt_rng.h"
#include "pt_scene.h"

namespace pt {

struct PrimHit {
    float t;
    V3 n;     // local-space normal
    int face; // box: index of the side that was hit (order of box::box, primitive.h:233-240); else 0
};

struct Hit {
    int32_t inst; // -1: miss
    float t;
    V3 p, n;
    float u, v;
    int face; // box: the side that was hit (see PrimHit); kept so that u, v can be recomputed later (hit_uv)
};

PT_HD V3 shuffle(V3 v, int plane)
{
    // primitive.h:104-121
    if (plane == B200PT_PLANE_XY)
        return v3(v.x, v.z, v.y);
    if (plane == B200PT_PLANE_YZ)
        return v3(v.y, v.x, v.z);
    return v;
}

PT_HD bool hit_sphere(V3 center, float radius, V3 o, V3 d, float t_min, float t_max, PrimHit &h)
{
    V3 oc = o - center;
    float a = dot(d, d);
    float b = dot(oc, d);
    float c = dot(oc, oc) - radius * radius;
    float discriminant = b * b - a * c;
    if (discriminant > 0) {
        float sq = sqrtf(discriminant);
        float temp = (-b - sq) / a;
        if (!(temp < t_max && temp > t_min)) {
            temp = (-b + sq) / a;
            if (!(temp < t_max && temp > t_min))
                return false;
        }
        h.t = temp;
        V3 p = o + temp * d;
        h.n = (p - center) / radius;
        h.face = 0;
        return true;
    }
    return false;
}

PT_HD bool hit_rect(int plane, bool flip, float x0, float z0, float x1, float z1, float y, V3 o, V3 d, float t0,
                    float t1, PrimHit &h)
{
    V3 to = shuffle(o, plane);
    V3 td = shuffle(d, plane);
    float t = fdiv(y - to.y, td.y);
    if (t < t0 || t > t1) // NB a NaN t passes, as in the reference
        return false;
    float xh = to.x + t * td.x;
    float zh = to.z + t * td.z;
    if (xh < x0 || xh > x1 || zh < z0 || zh > z1)
        return false;
    h.t = t; // rec.u / rec.v (primitive.h:206-207) are evaluated for the closest hit only: rect_uv()
    h.face = 0;
    V3 n = shuffle(v3(0.f, flip ? -1.f : 1.f, 0.f), plane);
    if (dot(d, n) > 0) // two_sided is hard-wired true (primitive.h:181,214-222)
        n = -n;
    h.n = n;
    return true;
}

// rec.u, rec.v of a rect hit (primitive.h:206-207), from the same expressions as the hit test
PT_HD void rect_uv(int plane, float x0, float z0, float x1, float z1, V3 o, V3 d, float t, float &u, float &v)
{
    V3 to = shuffle(o, plane);
    V3 td = shuffle(d, plane);
    float xh = to.x + t * td.x;
    float zh = to.z + t * td.z;
    u = (xh - x0) / (x1 - x0);
    v = (zh - x0) / (z1 - z0); // sic: x0 (primitive.h:207)
}

// box::hit = hittable_list::hit over six rects (primitive.h:233-246, hittable_list.h:21-38): sides in the order
// z = p0.z, z = p1.z, x = p0.x, x = p1.x, y = p0.y, y = p1.y; each is tested against [t_min, closest so far] with an
// INCLUSIVE upper bound (rect::hit), so a later side at the same t replaces an earlier one.  One loop body serves
// the three axis pairs by rotating (a, b, c) <- (b, c, a); every t and bound test is the rect's own expression.
PT_HD bool hit_box(const float *p, V3 o, V3 d, float t_min, float t_max, PrimHit &h)
{
    float oa = o.z, ob = o.x, oc = o.y, da = d.z, db = d.x, dc = d.y;
    float la = p[2], lb = p[0], lc = p[1], ha = p[5], hb = p[3], hc = p[4];
    float closest = t_max;
    int face = -1;
    float nsign = 0.f;
#if defined(__CUDA_ARCH__)
#pragma unroll // rolled, the rotation below is 12 moves per round, executed by the few lanes that reach a box
#endif
    for (int it = 0; it < 3; it++) {
        float t = fdiv(la - oa, da);
        if (!(t < t_min || t > closest)) { // NB a NaN t passes, as in the reference
            float bh = ob + t * db, ch = oc + t * dc;
            if (!(bh < lb || bh > hb || ch < lc || ch > hc)) {
                closest = t;
                face = 2 * it;
                nsign = da < 0 ? 1.f : -1.f; // normal -e_a, turned to face the ray (primitive.h:214-222)
            }
        }
        t = fdiv(ha - oa, da);
        if (!(t < t_min || t > closest)) {
            float bh = ob + t * db, ch = oc + t * dc;
            if (!(bh < lb || bh > hb || ch < lc || ch > hc)) {
                closest = t;
                face = 2 * it + 1;
                nsign = da > 0 ? -1.f : 1.f;
            }
        }
        float r;
        r = oa, oa = ob, ob = oc, oc = r;
        r = da, da = db, db = dc, dc = r;
        r = la, la = lb, lb = lc, lc = r;
        r = ha, ha = hb, hb = hc, hc = r;
    }
    if (face < 0)
        return false;
    h.t = closest;
    h.face = face;
    h.n = face < 2 ? v3(0.f, 0.f, nsign) : (face < 4 ? v3(nsign, 0.f, 0.f) : v3(0.f, nsign, 0.f));
    return true;
}

// The two calls constant_medium::hit makes on a box boundary (volume.h:39-41): box::hit over (-inf, inf) -> ta, then
// box::hit over (ta + 0.0001, inf) -> tb.  A side's t and its bound test do not depend on the range, so the six
// sides are evaluated once and the list walk of hit_box (inclusive upper bound: a later side at the same t
// replaces an earlier one) is replayed twice over the stored t's; only t is used by the medium.
PT_HD bool box_entry_exit(const float *p, V3 o, V3 d, float &ta, float &tb)
{
    float oa = o.z, ob = o.x, oc = o.y, da = d.z, db = d.x, dc = d.y;
    float la = p[2], lb = p[0], lc = p[1], ha = p[5], hb = p[3], hc = p[4];
    float ts[6];
    bool in[6];
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int it = 0; it < 3; it++) {
        float t = fdiv(la - oa, da);
        float bh = ob + t * db, ch = oc + t * dc;
        ts[2 * it] = t;
        in[2 * it] = !(bh < lb || bh > hb || ch < lc || ch > hc);
        t = fdiv(ha - oa, da);
        bh = ob + t * db, ch = oc + t * dc;
        ts[2 * it + 1] = t;
        in[2 * it + 1] = !(bh < lb || bh > hb || ch < lc || ch > hc);
        float r;
        r = oa, oa = ob, ob = oc, oc = r;
        r = da, da = db, db = dc, dc = r;
        r = la, la = lb, lb = lc, lc = r;
        r = ha, ha = hb, hb = hc, hc = r;
    }
    float closest = PT_FLT_MAX;
    bool any = false;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int k = 0; k < 6; k++)
        if (!(ts[k] < -PT_FLT_MAX || ts[k] > closest) && in[k]) { // NB a NaN t passes, as in rect::hit
            closest = ts[k];
            any = true;
        }
    if (!any)
        return false;
    ta = closest;
    const float t_min = (float)((double)ta + 0.0001); // `rec1.t + 0.0001` is a double sum
    closest = PT_FLT_MAX;
    any = false;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int k = 0; k < 6; k++)
        if (!(ts[k] < t_min || ts[k] > closest) && in[k]) {
            closest = ts[k];
            any = true;
        }
    tb = closest;
    return any;
}

// the bare primitive (sphere | rect | box) in its local frame
PT_HD bool hit_prim(uint32_t meta, const float *p, V3 o, V3 d, float t_min, float t_max, PrimHit &h)
{
    int kind = meta_kind(meta);
    if (kind == B200PT_PRIM_SPHERE)
        return hit_sphere(v3(p[0], p[1], p[2]), p[3], o, d, t_min, t_max, h);
    if (kind == B200PT_PRIM_RECT)
        return hit_rect(meta_plane(meta), meta_flip(meta), p[0], p[1], p[2], p[3], p[4], o, d, t_min, t_max, h);
    return hit_box(p, o, d, t_min, t_max, h);
}

// Where a traversal gets the medium variates from.  One Philox call serves two medium slots ({xi, xi2} each), and
// the last block drawn is kept: a ray that meets the same media again (the usual case: a scene-wide mist) pays for
// the generator once.
struct MediumRng {
    RngKey key;
    uint32_t stage; // stage word; counter index = medium slot / 2
    float xi;       // >= 0: explicit variate for every medium (b200pt_trace_batch); < 0: use the Philox stream
    mutable int32_t cached_block = -1;
    mutable F4 cached;
};

// the variates of medium `slot`: xi for volume.h:70, xi2 for the second copy of a BVH-duplicated medium
PT_HD void medium_variates(const MediumRng &mr, uint32_t slot, bool dup, float &xi, float &xi2)
{
    xi2 = -1.f;
    if (mr.xi >= 0.f) {
        xi = mr.xi;
        return;
    }
    const int32_t block = (int32_t)(slot >> 1);
    if (block != mr.cached_block) {
        mr.cached = rng4(mr.key, mr.stage, (uint32_t)block);
        mr.cached_block = block;
    }
    const bool odd = slot & 1u;
    xi = odd ? mr.cached.z : mr.cached.x;
    if (dup)
        xi2 = odd ? mr.cached.w : mr.cached.y;
}

// first root of the sphere quadratic inside (t_lo, t_hi), in sphere::hit's order (primitive.h:72-92)
PT_HD bool sphere_pick(float r1, float r2, float t_lo, float t_hi, float &t)
{
    if (r1 < t_hi && r1 > t_lo) {
        t = r1;
        return true;
    }
    if (r2 < t_hi && r2 > t_lo) {
        t = r2;
        return true;
    }
    return false;
}

// constant_medium::hit (volume.h:29-93): boundary->hit over (-inf, inf) gives the entry, boundary->hit from
// entry + 0.0001 the exit; the event distance is drawn only when the clipped interval is not empty (volume.h:70).
// A BVH-duplicated medium (bvh.h:150-153) is intersected twice by the reference and the nearer event wins: the
// minimum of two free-flight draws.
PT_HD bool hit_medium(uint32_t meta, const float *p, float density, V3 o, V3 d, float t_min, float t_max,
                      const MediumRng &mr, PrimHit &h)
{
    float ta, tb;
    if (meta_kind(meta) == B200PT_PRIM_SPHERE) {
        // both calls of sphere::hit evaluate the same two roots
        V3 center = v3(p[0], p[1], p[2]);
        V3 oc = o - center;
        float a = dot(d, d);
        float b = dot(oc, d);
        float c = dot(oc, oc) - p[3] * p[3];
        float discriminant = b * b - a * c;
        if (!(discriminant > 0))
            return false;
        float sq = sqrtf(discriminant);
        float r1 = (-b - sq) / a, r2 = (-b + sq) / a;
        if (!sphere_pick(r1, r2, -PT_FLT_MAX, PT_FLT_MAX, ta))
            return false;
        if (!sphere_pick(r1, r2, (float)((double)ta + 0.0001), PT_FLT_MAX, tb)) // `rec1.t + 0.0001` is a double sum
            return false;
    } else if (meta_kind(meta) == B200PT_PRIM_BOX) {
        if (!box_entry_exit(p, o, d, ta, tb))
            return false;
    } else {
        // a rect boundary: the second boundary->hit starts 0.0001 behind the only crossing and finds nothing; a ray
        // IN the plane gives t = NaN twice, and every comparison below is then false up to `hit_distance < NaN`
        return false;
    }
    if (ta < t_min)
        ta = t_min;
    if (tb > t_max)
        tb = t_max;
    if (ta >= tb)
        return false;
    if (ta < 0)
        ta = 0;
    float xi, xi2;
    medium_variates(mr, meta_medium_slot(meta), meta_dup(meta), xi, xi2);
    float dlen = length(d);
    float distance_inside_boundary = (tb - ta) * dlen;
    float hit_distance = -(1.f / density) * logf(xi);
    if (xi2 >= 0.f) {
        float hd2 = -(1.f / density) * logf(xi2);
        hit_distance = hd2 < hit_distance ? hd2 : hit_distance;
    }
    if (hit_distance < distance_inside_boundary) {
        h.t = ta + hit_distance / dlen;
        h.n = v3(1.f, 0.f, 0.f); // "arbitrary" (volume.h:86)
        h.face = 0;
        return true;
    }
    return false;
}

struct InstanceData { // one PtInstance in registers
    float inv[12];
    uint32_t meta;
    float p[6];
    float density;
};

PT_HD void load_instance(const PtInstance *table, int idx, InstanceData &I)
{
#if defined(__CUDA_ARCH__)
    const float4 *q = reinterpret_cast<const float4 *>(table + idx);
    // plain loads: `table` is in global memory, or in shared memory for small scenes (kernels.cu SmallScene)
    float4 a = q[0], b = q[1], c = q[2], d = q[3], e = q[4];
    I.inv[0] = a.x, I.inv[1] = a.y, I.inv[2] = a.z, I.inv[3] = a.w;
    I.inv[4] = b.x, I.inv[5] = b.y, I.inv[6] = b.z, I.inv[7] = b.w;
    I.inv[8] = c.x, I.inv[9] = c.y, I.inv[10] = c.z, I.inv[11] = c.w;
    I.meta = __float_as_uint(d.x);
    I.p[0] = d.y, I.p[1] = d.z, I.p[2] = d.w;
    I.p[3] = e.x, I.p[4] = e.y, I.p[5] = e.z;
    I.density = e.w;
#else
    const PtInstance &s = table[idx];
    for (int i = 0; i < 12; i++)
        I.inv[i] = s.inv[i];
    I.meta = s.meta;
    for (int i = 0; i < 6; i++)
        I.p[i] = s.p[i];
    I.density = s.density;
#endif
}

// instance::hit without the world-space epilogue: local ray, local hit record, t preserved (primitive.h:298-301).
// MEDIA = false compiles the constant-medium branch out (scenes without volumes: smaller kernels).
template <bool MEDIA>
PT_HD bool hit_instance(const InstanceData &I, V3 o, V3 d, float t_min, float t_max, const MediumRng &mr, PrimHit &h)
{
    V3 lo = xform_point(I.inv, o);
    V3 ld = xform_vector(I.inv, d);
    if (!MEDIA || !meta_medium(I.meta))
        return hit_prim(I.meta, I.p, lo, ld, t_min, t_max, h);
    return hit_medium(I.meta, I.p, I.density, lo, ld, t_min, t_max, mr, h);
}

// aabb::hit (aabb.h:34-53) of one child half, used as a culling test against [t_min, t_limit] (inclusive).
// (b - o) * (1/d) is the reference's own expression; a primitive's t comes from a true division, so the two can
// differ by an ulp: the caller passes t_limit = closest t so far widened by a few ulps, which can only let a ray
// visit a subtree it did not need, never skip one whose primitive it hits -- including a primitive at EXACTLY
// the current closest t, which must still be offered to the tie rule in trace_closest.
PT_HD bool hit_aabb(const float *lo, const float *hi, V3 o, V3 inv_d, float t_min, float t_limit, float &t_entry)
{
    float t0 = (lo[0] - o.x) * inv_d.x, t1 = (hi[0] - o.x) * inv_d.x;
    float tn = fminf(t0, t1), tf = fmaxf(t0, t1);
    t_min = fmaxf(tn, t_min); // fmaxf/fminf drop NaNs like the reference's `t0 > tmin ? t0 : tmin`
    t_limit = fminf(tf, t_limit);
    t0 = (lo[1] - o.y) * inv_d.y, t1 = (hi[1] - o.y) * inv_d.y;
    tn = fminf(t0, t1), tf = fmaxf(t0, t1);
    t_min = fmaxf(tn, t_min);
    t_limit = fminf(tf, t_limit);
    t0 = (lo[2] - o.z) * inv_d.z, t1 = (hi[2] - o.z) * inv_d.z;
    tn = fminf(t0, t1), tf = fmaxf(t0, t1);
    t_min = fmaxf(tn, t_min);
    t_limit = fminf(tf, t_limit);
    t_entry = t_min;
    return t_limit >= t_min;
}

// closest t so far, widened by ~3 ulps for the culling tests
PT_HD float cull_limit(float best_t) { return best_t + fabsf(best_t) * 4e-7f; }

// rec.u / rec.v: only rect::hit sets them (primitive.h:206-207; a box side is a rect); sphere and constant_medium
// leave them as they were (Q11) -- reported as 0.  (lo, ld) = the ray in the instance's frame.
template <bool MEDIA> PT_HD void local_uv(const InstanceData &I, V3 lo, V3 ld, float t, int face, float &u, float &v)
{
    u = v = 0.f;
    if (MEDIA && meta_medium(I.meta))
        return;
    const int kind = meta_kind(I.meta);
    const float *p = I.p;
    if (kind == B200PT_PRIM_RECT) {
        rect_uv(meta_plane(I.meta), p[0], p[1], p[2], p[3], lo, ld, t, u, v);
    } else if (kind == B200PT_PRIM_BOX) {
        const int f = face >> 1;
        if (f == 0)
            rect_uv(B200PT_PLANE_XY, p[0], p[1], p[3], p[4], lo, ld, t, u, v);
        else if (f == 1)
            rect_uv(B200PT_PLANE_YZ, p[1], p[2], p[4], p[5], lo, ld, t, u, v);
        else
            rect_uv(B200PT_PLANE_XZ, p[0], p[2], p[3], p[5], lo, ld, t, u, v);
    }
}

// the same u, v from what the wavefront keeps of a hit (the ray, t, the instance and the box side): the shade
// kernels call this only when the scene has image textures
PT_HD void hit_uv(const PtSceneView &s, int inst, int face, V3 o, V3 d, float t, float &u, float &v)
{
    InstanceData I;
    load_instance(s.inst, inst, I);
    local_uv<true>(I, xform_point(I.inv, o), xform_vector(I.inv, d), t, face, u, v);
}

// -DPT_DEBUG_BOUNDS: range checks of every node, instance and stack index on the device (tools/variants_smoke.py)
#if defined(PT_DEBUG_BOUNDS) && defined(__CUDA_ARCH__)
#define PT_CHECK(cond)                                                                                                 \
    do {                                                                                                               \
        if (!(cond))                                                                                                   \
            __trap();                                                                                                  \
    } while (0)
#else
#define PT_CHECK(cond) ((void)0)
#endif

struct TraceCounters {
    uint32_t boxes, prims;
};

#define PT_MAX_DIRECT_EMITTERS 8 // more emitters than this: shadow rays fall back to the plain closest-hit query

// World::hit (world.h:17-20) as a resumable state machine: init() -> steps until one returns false -> result().
// A step visits one inner node (node_step) or tests one leaf (leaf_step).  The CPU restatement runs a ray to
// completion (run / trace_closest); the kernels run 32 rays per warp and vote on the kind of step (run_warp).
struct TraversalStack {
    int32_t ref[PT_STACK_DEPTH];
    float t[PT_STACK_DEPTH];
};

struct Traversal {
    V3 o, d, inv_d;
    float t_min, best_t;
    int32_t best_inst;
    PrimHit best;
    int32_t cur; // node to visit next: >= 0 inner node, < 0 leaf (~instance)
    int32_t sp;
    // The pending-node stack is an object of its own (attach it before init): inside this struct, every store to
    // stack_ref[sp] might have overwritten sp or the ray as far as the compiler could tell, and the whole state was
    // kept in local memory and re-read after each push.
    TraversalStack *stk;

    PT_HD void init(V3 origin, V3 dir, float tmin, float tmax)
    {
        o = origin;
        d = dir;
        inv_d = v3(1.0f / dir.x, 1.0f / dir.y, 1.0f / dir.z);
        t_min = tmin;
        best_t = tmax;
        best_inst = -1;
        best.t = 0.f;
        best.n = v3(0.f, 0.f, 0.f);
        best.face = 0;
        cur = 0; // root
        sp = 0;
    }

    // ANY (occlusion walk): best_t does not shrink before the walk ends, so an entry that passed the slab test
    // when it was pushed passes it now -- its entry distance is neither stored nor compared.
    template <bool ANY = false> PT_HD bool pop()
    {
        if (ANY) {
            if (sp > 0) {
                cur = stk->ref[--sp];
                return true;
            }
            return false;
        }
        const float lim = cull_limit(best_t);
        while (sp > 0) {
            sp--;
            if (stk->t[sp] <= lim) {
                cur = stk->ref[sp];
                return true;
            }
        }
        return false;
    }

    // One inner node (cur >= 0) of the 4-wide tree: slab tests of its (up to) four children, descend into the nearest
    // one that is hit, push the others farthest first.  Returns false when the traversal is complete.
    template <bool COUNT, bool SORT = true> PT_HD bool node_step(const PtSceneView &s, TraceCounters *counters)
    {
        PT_CHECK(cur >= 0 && cur < s.n_nodes);
        const PtNode4 *node = s.nodes + cur;
        float e[4];
        int32_t ref[4];
#if defined(__CUDA_ARCH__)
        const float4 *q = reinterpret_cast<const float4 *>(node);
        // (loading the node with the L1 evict-last priority, ld.global.nc.L1::evict_last, changes nothing: measured in round 2)
        const float4 lx = __ldg(q), ly = __ldg(q + 1), lz = __ldg(q + 2), hx = __ldg(q + 3), hy = __ldg(q + 4),
                     hz = __ldg(q + 5);
        const int4 r = __ldg(reinterpret_cast<const int4 *>(q + 6));
        const float lox[4] = {lx.x, lx.y, lx.z, lx.w}, loy[4] = {ly.x, ly.y, ly.z, ly.w}, loz[4] = {lz.x, lz.y, lz.z, lz.w};
        const float hix[4] = {hx.x, hx.y, hx.z, hx.w}, hiy[4] = {hy.x, hy.y, hy.z, hy.w}, hiz[4] = {hz.x, hz.y, hz.z, hz.w};
        ref[0] = r.x, ref[1] = r.y, ref[2] = r.z, ref[3] = r.w;
#else
        const float *lox = node->lox, *loy = node->loy, *loz = node->loz, *hix = node->hix, *hiy = node->hiy,
                    *hiz = node->hiz;
        for (int k = 0; k < 4; k++)
            ref[k] = node->ref[k];
#endif
        const float lim = cull_limit(best_t);
        const float never = __builtin_huge_valf();
        bool in[4];
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
        for (int k = 0; k < 4; k++) {
            const float lo[3] = {lox[k], loy[k], loz[k]}, hi[3] = {hix[k], hiy[k], hiz[k]};
            const bool used = ref[k] != PT_NO_CHILD;
            if (COUNT && used)
                counters->boxes += 1;
            // no short circuit: the slab test of an empty slot (a zero box) is computed and discarded -- a branch
            // around it cost five instructions per child, and empty slots are rare
            in[k] = hit_aabb(lo, hi, o, inv_d, t_min, lim, e[k]) && used;
        }
        // scene_pack refuses a tree whose traversal could need more than PT_STACK_DEPTH pending entries
        // (bvh_build.cpp collapse_node computes the bound), so the pushes below are not range-checked
        if (!SORT) {
            // occlusion walk: any order will do and nothing is compared on the way out of the stack.  Children are
            // visited in slot order: the first one that is hit next, the others pushed last slot first -- no sorting
            // network, three predicated stores.  (On the 8192-instance scene, oracle/walk_stats.cpp: 27.0 slab + 1.65
            // primitive tests per shadow ray, against 27.7 + 1.60 with the nearest child first and 31.6 + 1.41 in the
            // opposite slot order.)
            if (in[3] && (in[0] || in[1] || in[2])) {
                PT_CHECK(sp < PT_STACK_DEPTH);
                stk->ref[sp++] = ref[3];
            }
            if (in[2] && (in[0] || in[1])) {
                PT_CHECK(sp < PT_STACK_DEPTH);
                stk->ref[sp++] = ref[2];
            }
            if (in[1] && in[0]) {
                PT_CHECK(sp < PT_STACK_DEPTH);
                stk->ref[sp++] = ref[1];
            }
            if (!(in[0] || in[1] || in[2] || in[3]))
                return pop<true>();
            cur = in[0] ? ref[0] : (in[1] ? ref[1] : (in[2] ? ref[2] : ref[3]));
            return true;
        }
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
        for (int k = 0; k < 4; k++)
            e[k] = in[k] ? e[k] : never;
        // sorting network, nearest first.  The distances move through min / max and the references through two selects
        // (an entry distance is never NaN: hit_aabb's maxima drop NaNs) -- five instructions per exchange
#define PT_CSWAP(a, b)                                                                                                 \
    {                                                                                                                  \
        const bool sw = e[b] < e[a];                                                                                   \
        const float e_lo = fminf(e[a], e[b]), e_hi = fmaxf(e[a], e[b]);                                                \
        const int32_t r_lo = sw ? ref[b] : ref[a], r_hi = sw ? ref[a] : ref[b];                                        \
        e[a] = e_lo, e[b] = e_hi, ref[a] = r_lo, ref[b] = r_hi;                                                        \
    }
        PT_CSWAP(0, 1)
        PT_CSWAP(2, 3)
        PT_CSWAP(0, 2)
        PT_CSWAP(1, 3)
        PT_CSWAP(1, 2)
#undef PT_CSWAP
        if (!(e[0] < never))
            return pop<false>();
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
        for (int k = 3; k >= 1; k--) {
            PT_CHECK(!(e[k] < never) || sp < PT_STACK_DEPTH);
            if (e[k] < never) {
                stk->ref[sp] = ref[k];
                stk->t[sp] = e[k];
                sp++;
            }
        }
        cur = ref[0];
        return true;
    }

    // One leaf: instance idx against [t_min, best_t].  Returns true when the traversal must end here.
    // ANY = true is the occlusion phase of a shadow query (see emitters_first): emitters were tested already and the
    // first accepted hit ends the traversal.
    template <bool COUNT, bool MEDIA, bool ANY = false>
    PT_HD bool leaf_test(const PtSceneView &s, int idx, const MediumRng &mr, TraceCounters *counters)
    {
        PT_CHECK(idx >= 0 && idx < s.n_inst);
        InstanceData I;
        load_instance(s.inst, idx, I);
        if (ANY && meta_emitter(I.meta))
            return false;
        PrimHit h;
        if (COUNT)
            counters->prims += 1;
        // Every primitive test is inclusive at its upper bound or never ties (rect: `t > t1` rejects,
        // primitive.h:196; sphere: strict), so a surface at exactly best_t is reported again.  The reference hands
        // the unshrunk t_max to both children and keeps the RIGHT record unless left.t < right.t (bvh.h:39-48):
        // among equal t the right-most leaf of ITS tree wins.  The instance table is stored in that leaf order
        // (scene_compile.cpp), hence: a tie goes to the higher index.
        if (hit_instance<MEDIA>(I, o, d, t_min, best_t, mr, h) && (h.t < best_t || idx > best_inst)) {
            best_t = h.t;
            best = h;
            best_inst = idx;
            if (ANY)
                return true; // occluded: nothing nearer can change the answer
        }
        return false;
    }

    // One leaf step (cur < 0): the instance ~cur, then the next pending node.  Returns false when the traversal is
    // complete.
    template <bool COUNT, bool MEDIA, bool ANY = false>
    PT_HD bool leaf_step(const PtSceneView &s, const MediumRng &mr, TraceCounters *counters)
    {
        if (leaf_test<COUNT, MEDIA, ANY>(s, ~cur, mr, counters))
            return false;
        return pop<ANY>();
    }

    // Shadow queries (integrator.h:246-252) need only the emission of whatever the ray meets first, and that is zero
    // unless the first thing is an emitter.  So: the few emitters are intersected directly (closest one wins, same
    // tie rule), then the BVH is walked for ANY other instance in front of it.  Same answer as the closest-hit
    // query whenever that answer matters; a ray that meets no emitter is done without touching the BVH.
    // Returns whether an emitter was hit (best_* hold it).  Requires s.n_emitters <= PT_MAX_DIRECT_EMITTERS.
    // known_inst >= 0: that emitter's hit over (t_min, FLT_MAX) is already known (NEE has just intersected the light
    // it sampled to get its pdf, pt_shade.h light_pdf_value): `known` is taken as is and the instance is not
    // intersected again.
    template <bool COUNT>
    PT_HD bool emitters_first(const PtSceneView &s, TraceCounters *counters, int known_inst = -1,
                              const PrimHit *known = nullptr)
    {
        MediumRng none;
        none.xi = -1.f;
        if (known_inst >= 0) {
            best_t = known->t;
            best = *known;
            best_inst = known_inst;
        }
        for (int e = 0; e < s.n_emitters; e++) {
            const int idx = s.emitters[e];
            if (idx == known_inst)
                continue;
            InstanceData I;
            load_instance(s.inst, idx, I);
            PrimHit h;
            if (COUNT)
                counters->prims += 1;
            if (hit_instance<false>(I, o, d, t_min, best_t, none, h) && (h.t < best_t || idx > best_inst)) {
                best_t = h.t;
                best = h;
                best_inst = idx;
            }
        }
        return best_inst >= 0;
    }

    // Small scenes: no tree.  All instances in index order; a rect (most of a Cornell box) is tested directly --
    // its t needs one row of the transform, the rest is computed only when t is in range -- anything else behind
    // the slab test of its leaf box, exactly the test the tree would have made.  The closest hit is the same:
    // every primitive test depends on (ray, t_min, best_t) only and ties go to the higher index either way.
    // On the GPU the 32 lanes walk the table together (tab in shared memory): no stack, no divergence between
    // primitive kinds.
    template <bool COUNT, bool MEDIA, bool ANY = false>
    PT_HD void run_small(const PtBox *boxes, const PtInstance *table, int n, const MediumRng &mr,
                         TraceCounters *counters)
    {
        for (int idx = 0; idx < n; idx++) {
            const PtInstance &E = table[idx];
            const uint32_t meta = E.meta;
            if (ANY && meta_emitter(meta))
                continue;
            PrimHit h;
            bool ok;
            if (meta_kind(meta) == B200PT_PRIM_RECT && !(MEDIA && meta_medium(meta))) {
                if (COUNT)
                    counters->prims += 1;
                // shuffle() puts the plane's axis in .y: XY -> local z, YZ -> local x, XZ -> local y
                const int plane = meta_plane(meta);
                const float *row = E.inv + (plane == B200PT_PLANE_XY ? 8 : (plane == B200PT_PLANE_YZ ? 0 : 4));
                const float oy = row[0] * o.x + row[1] * o.y + row[2] * o.z + row[3]; // xform_point, that row
                const float dy = row[0] * d.x + row[1] * d.y + row[2] * d.z;          // xform_vector, that row
                const float t = fdiv(E.p[4] - oy, dy);
                if (t < t_min || t > best_t) // rect::hit's own range test (a NaN t passes)
                    continue;
                if (ANY) {
                    // occlusion walks rarely get here, and their kernel is faster with the shared routine (smaller code)
                    ok = hit_rect(plane, meta_flip(meta), E.p[0], E.p[1], E.p[2], E.p[3], E.p[4],
                                  xform_point(E.inv, o), xform_vector(E.inv, d), t_min, best_t, h);
                } else {
                    // the rest of rect::hit (primitive.h:196-222) from the two other rows, each the expression
                    // xform_point / xform_vector evaluate: hit_rect() would redo the third row and the division
                    const float *rx = E.inv + (plane == B200PT_PLANE_YZ ? 4 : 0);
                    const float *rz = E.inv + (plane == B200PT_PLANE_XY ? 4 : 8);
                    const float ox = rx[0] * o.x + rx[1] * o.y + rx[2] * o.z + rx[3];
                    const float dx = rx[0] * d.x + rx[1] * d.y + rx[2] * d.z;
                    const float oz = rz[0] * o.x + rz[1] * o.y + rz[2] * o.z + rz[3];
                    const float dz = rz[0] * d.x + rz[1] * d.y + rz[2] * d.z;
                    const float xh = ox + t * dx, zh = oz + t * dz;
                    if (xh < E.p[0] || xh > E.p[2] || zh < E.p[1] || zh > E.p[3])
                        continue;
                    const float sgn = meta_flip(meta) ? -1.f : 1.f;
                    V3 n = plane == B200PT_PLANE_XY ? v3(0.f, 0.f, sgn)
                                                    : (plane == B200PT_PLANE_YZ ? v3(sgn, 0.f, 0.f) : v3(0.f, sgn, 0.f));
                    // dot(local d, n): the two products with the zero components only matter when they are NaN
                    if ((dx * 0.f + dz * 0.f) + dy * sgn > 0)
                        n = -n;
                    h.t = t;
                    h.face = 0;
                    h.n = n;
                    ok = true;
                }
            } else {
                float entry;
                if (COUNT)
                    counters->boxes += 1;
                if (!hit_aabb(boxes[idx].lo, boxes[idx].hi, o, inv_d, t_min, cull_limit(best_t), entry))
                    continue;
                if (COUNT)
                    counters->prims += 1;
                InstanceData I;
                load_instance(table, idx, I);
                ok = hit_instance<MEDIA>(I, o, d, t_min, best_t, mr, h);
            }
            if (ok && (h.t < best_t || idx > best_inst)) {
                best_t = h.t;
                best = h;
                best_inst = idx;
                if (ANY)
                    return;
            }
        }
    }

    // one ray run to completion (CPU restatement, and any caller without a warp to vote with)
    template <bool COUNT, bool MEDIA, bool ANY = false>
    PT_HD void run(const PtSceneView &s, const MediumRng &mr, TraceCounters *counters)
    {
        bool more = true;
        while (more)
            more = cur >= 0 ? node_step<COUNT, !ANY>(s, counters) : leaf_step<COUNT, MEDIA, ANY>(s, mr, counters);
    }

#if defined(__CUDACC__)
    // The same state machine driven by the whole warp: every round the warp votes and runs ONE kind of step --
    // the one more of its lanes are waiting for.  Rays need different numbers of node visits between two leaves;
    // letting each lane loop on its own (while-while) left 14 of 32 lanes busy in the node loop and 6 in the leaf
    // code on the 8192-instance scene (ncu, profiles/r1b_*).  With the vote the step that runs always has at least
    // half of the unfinished lanes in it, and the code of the other kind is not interleaved with it.
    // Must be called by all 32 lanes (`active` = this lane has a ray).  Per ray the sequence of steps is the one
    // run() takes, so the result is identical.
    template <bool COUNT, bool MEDIA, bool ANY = false>
    __device__ __forceinline__ void run_warp(const PtSceneView &s, bool active, const MediumRng &mr, TraceCounters *counters)
    {
#ifdef PT_LANE_LOOP // A/B: every lane loops on its own
        if (active)
            run<COUNT, MEDIA, ANY>(s, mr, counters);
        return;
#endif
        while (true) {
            const unsigned m_act = __ballot_sync(0xffffffffu, active);
            if (!m_act)
                break;
            const unsigned m_in = __ballot_sync(0xffffffffu, active && cur >= 0);
            if (2 * __popc(m_in) >= __popc(m_act)) {
                if (active && cur >= 0)
                    active = node_step<COUNT, !ANY>(s, counters);
            } else {
                if (active && cur < 0)
                    active = leaf_step<COUNT, MEDIA, ANY>(s, mr, counters);
            }
        }
    }
#endif

#if defined(__CUDACC__)
    // while-while: inner nodes until no lane of the warp wants one any more, then the leaves that were reached,
    // together.  Same per-ray sequence of steps as run().
    // DIV > 0: the node phase ends as soon as fewer than 1 / DIV of the unfinished lanes still want an inner node -- the
    // many lanes waiting at their leaves are served first and the few stragglers walk on together with the lanes that
    // pop their next node after the leaf (DIV = 2 is run_warp's vote, DIV = 0 the plain while-while).
    template <bool COUNT, bool MEDIA, bool ANY = false, int DIV = 0>
    __device__ __forceinline__ void run_warp_nodes_first(const PtSceneView &s, bool active, const MediumRng &mr,
                                                         TraceCounters *counters)
    {
        if (DIV == 0) {
            while (__any_sync(0xffffffffu, active)) {
                while (__any_sync(0xffffffffu, active && cur >= 0))
                    if (active && cur >= 0)
                        active = node_step<COUNT, !ANY>(s, counters); // an occlusion query does not care for the order
                if (active) // cur < 0 here
                    active = leaf_step<COUNT, MEDIA, ANY>(s, mr, counters);
            }
            return;
        }
        while (true) {
            unsigned m_act = __ballot_sync(0xffffffffu, active);
            if (!m_act)
                break;
            while (true) {
                const unsigned m_in = __ballot_sync(0xffffffffu, active && cur >= 0);
                if (__popc(m_in) * DIV < __popc(m_act) || !m_in)
                    break;
                if (active && cur >= 0)
                    active = node_step<COUNT, !ANY>(s, counters);
                m_act = __ballot_sync(0xffffffffu, active);
            }
            if (active && cur < 0)
                active = leaf_step<COUNT, MEDIA, ANY>(s, mr, counters);
        }
    }
#endif

    // instance::hit epilogue (primitive.h:302-305): p and normal back to world space, rec.u / rec.v
    // want_p = false (shadow hits on emitters with constant textures): only instance, t and normal are filled in;
    // want_n = false (... that are all two-sided): not even the normal -- emission does not look at it
    template <bool MEDIA> PT_HD bool result(const PtSceneView &s, Hit &out, bool want_p = true, bool want_n = true) const
    {
        if (best_inst < 0) {
            out.inst = -1;
            return false;
        }
        if (!want_p && !want_n) {
            out.inst = best_inst;
            out.t = best.t;
            out.p = v3(0.f, 0.f, 0.f);
            out.n = v3(0.f, 0.f, 0.f);
            out.u = out.v = 0.f;
            out.face = best.face;
            return true;
        }
        InstanceData I;
        load_instance(s.inst, best_inst, I);
        if (!want_p) {
            out.inst = best_inst;
            out.t = best.t;
            out.p = v3(0.f, 0.f, 0.f);
            out.n = eigen_normalized(xform_vector_transposed(I.inv, best.n));
            out.u = out.v = 0.f;
            out.face = best.face;
            return true;
        }
        V3 lo = xform_point(I.inv, o);
        V3 ld = xform_vector(I.inv, d);
        V3 pl = lo + best.t * ld;
        const float *fwd = s.fwd[best_inst].m;
        out.inst = best_inst;
        out.t = best.t;
        out.p = xform_point(fwd, pl);
        out.n = eigen_normalized(xform_vector_transposed(I.inv, best.n));
        // rec.u / rec.v: only rect::hit sets them (primitive.h:206-207; a box side is a rect); sphere and
        // constant_medium leave them as they were (Q11) -- reported as 0
        out.face = best.face;
        local_uv<MEDIA>(I, lo, ld, best.t, best.face, out.u, out.v);
        return true;
    }
};

// World::hit (world.h:17-20): closest hit over the whole scene in [t_min, t_max], one ray run to completion.
template <bool COUNT, bool MEDIA = true>
PT_HD bool trace_closest(const PtSceneView &s, V3 o, V3 d, float t_min, float t_max, const MediumRng &mr, Hit &out,
                         TraceCounters *counters)
{
    Traversal T;
    TraversalStack T_stack;
    T.stk = &T_stack;
    T.init(o, d, t_min, t_max);
    if (s.n_small > 0)
        T.template run_small<COUNT, MEDIA>(s.small, s.inst, s.n_small, mr, counters);
    else
        T.template run<COUNT, MEDIA>(s, mr, counters);
    return T.template result<MEDIA>(s, out);
}

// The shadow query of NEE: true iff the FIRST thing the ray meets is an emitter (out = that hit).  A false return
// stands for "miss or a non-emitting surface first", which contribute the same nothing.
template <bool COUNT, bool MEDIA = true>
PT_HD bool trace_shadow(const PtSceneView &s, V3 o, V3 d, float t_min, float t_max, const MediumRng &mr, Hit &out,
                        TraceCounters *counters, int known_inst = -1, const PrimHit *known = nullptr)
{
    Traversal T;
    TraversalStack T_stack;
    T.stk = &T_stack;
    T.init(o, d, t_min, t_max);
    if (s.n_emitters > PT_MAX_DIRECT_EMITTERS) {
        T.template run<COUNT, MEDIA>(s, mr, counters);
        return T.template result<MEDIA>(s, out);
    }
    if (!T.template emitters_first<COUNT>(s, counters, known_inst, known))
        return false;
    const int32_t lit = T.best_inst;
    if (s.n_small > 0)
        T.template run_small<COUNT, MEDIA, true>(s.small, s.inst, s.n_small, mr, counters);
    else
        T.template run<COUNT, MEDIA, true>(s, mr, counters);
    if (T.best_inst != lit)
        return false;
    return T.template result<MEDIA>(s, out, s.emit_needs_p != 0, s.emit_needs_n != 0);
}

} // namespace pt
