// pt_shade.h -- what happens at a path vertex: textures, material scatter/emission, BSDF sampling and pdfs,
// light sampling and the next-event-estimation sample with its MIS weight.  One function per stage of the
// reference's NEEIterative::color loop (integrator.h:176-339) and RecursivePT::color (integrator.h:29-72); the
// quirks of the code at HEAD are kept on purpose and marked "sic".
#pragma once
#include "pt_intersect.h"

namespace pt {

struct RenderParams {
    int32_t width, height;
    int32_t integrator; // B200PT_INTEGRATOR_*
    int32_t max_bounces, light_samples;
    int32_t light_group; // light_samples rounded up to a power of two when <= 32 (k_nee's lane group), else 0
    int32_t russian_roulette, only_direct;
    int32_t corrected; // config.physically_corrected (pt_corrected.h)
    float normal_offset;
    uint32_t seed_lo, seed_hi;
};

// ---- textures (texture.h) ---------------------------------------------------------------------------------------

PT_HD_COLD float perlin_noise(const PtSceneView &s, V3 p)
{
    // perlin::noise (texture.h:132-156) + perlin_interp (texture.h:111-130): the smoothstep is applied twice (sic)
    float fx = floorf(p.x), fy = floorf(p.y), fz = floorf(p.z);
    float u = p.x - fx, v = p.y - fy, w = p.z - fz;
    u = u * u * (3 - 2 * u);
    v = v * v * (3 - 2 * v);
    w = w * w * (3 - 2 * w);
    int i = (int)fx, j = (int)fy, k = (int)fz;
    float uu = u * u * (3 - 2 * u);
    float vv = v * v * (3 - 2 * v);
    float ww = w * w * (3 - 2 * w);
    float accum = 0;
    for (int di = 0; di < 2; di++)
        for (int dj = 0; dj < 2; dj++)
            for (int dk = 0; dk < 2; dk++) {
                int idx = s.perlin_perm[(i + di) & 255] ^ s.perlin_perm[256 + ((j + dj) & 255)] ^
                          s.perlin_perm[512 + ((k + dk) & 255)];
                V3 c = v3(s.perlin_ranvec[idx * 3], s.perlin_ranvec[idx * 3 + 1], s.perlin_ranvec[idx * 3 + 2]);
                V3 weight_v = v3(u - di, v - dj, w - dk);
                accum += (di * uu + (1 - di) * (1 - uu)) * (dj * vv + (1 - dj) * (1 - vv)) *
                         (dk * ww + (1 - dk) * (1 - ww)) * dot(c, weight_v);
            }
    return accum;
}

PT_HD_COLD float checker_sines(float scale, V3 p) { return sinf(scale * p.x) * sinf(scale * p.y) * sinf(scale * p.z); }

// texture::value and texture::alpha in one walk (checker textures may nest, texture.h:46-68)
PT_HD void tex_eval(const PtSceneView &s, int id, float u, float v, V3 p, V3 &value, float &alpha)
{
#if defined(__CUDA_ARCH__)
#pragma unroll 1 // unrolled, the three sinf of a checker level are 1700 instructions in every kernel that shades
#endif
    for (int depth = 0; depth < 8; depth++) {
        const b200pt_texture &t = s.tex[id];
        if (t.kind == B200PT_TEX_IMAGE) {
            // image_texture::value / alpha (image.h:15-46): wrap by truncation, then nearest texel.  The reference
            // indexes past the row for u == 1 or v == 1 exactly; clamped here.
            // A NaN coordinate (a ray IN the rect's plane) is int(NaN) there, i.e. a crash; it reads texel 0 here.
            const PtImage &im = s.images[t.odd];
            u = is_nan(u) ? 0.f : u;
            v = is_nan(v) ? 0.f : v;
            v -= (float)(int)v;
            if (v < 0)
                v += 1;
            u -= (float)(int)u;
            if (u < 0)
                u += 1;
            int y = (int)(v * im.height), x = (int)(u * im.width);
            y = y < 0 ? 0 : (y >= im.height ? im.height - 1 : y);
            x = x < 0 ? 0 : (x >= im.width ? im.width - 1 : x);
            const PtTexel tx = s.texels[im.offset + (uint32_t)y * (uint32_t)im.width + (uint32_t)x];
            value = v3(tx.r, tx.g, tx.b);
            alpha = tx.a;
            return;
        }
        if (t.kind == B200PT_TEX_CHECKER) {
            float sines = checker_sines(t.scale, p);
            id = sines > 0 ? t.odd : t.even;
            continue;
        }
        if (t.kind == B200PT_TEX_PERLIN) {
            float nz = perlin_noise(s, t.scale * p);
            value = v3(nz, nz, nz); // vec3(1,1,1) * noise (texture.h:191); may be negative (sic)
            alpha = 1.0f;
            return;
        }
        value = v3(t.color[0], t.color[1], t.color[2]);
        alpha = t.alpha;
        return;
    }
    value = v3(0.f, 0.f, 0.f);
    alpha = 1.f;
}

// ---- materials (material.h) -------------------------------------------------------------------------------------

// material::emitted: only diffuse_light overrides it (material.h:209-228); isotropic's 3-argument `emitted`
// does not override the 5-argument virtual (material.h:263-266), so media never emit.
PT_HD V3 mat_emitted(const PtSceneView &s, const b200pt_material &m, V3 ray_dir, V3 n, float u, float v, V3 p)
{
    if (m.kind != B200PT_MAT_DIFFUSE_LIGHT)
        return v3(0.f, 0.f, 0.f);
    bool aligned = dot(n, ray_dir) > 0;
    if (!aligned || m.two_sided) {
        V3 value;
        float alpha;
        tex_eval(s, m.texture, u, v, p, value, alpha);
        return (m.power * value) * alpha;
    }
    return v3(0.f, 0.f, 0.f);
}

// material::scatter: returns did_scatter; `attenuation` is left untouched by a light (material.h:187-191), so the
// caller's value from the previous vertex survives (integrator.h:181,199).
PT_HD bool mat_scatter(const PtSceneView &s, const b200pt_material &m, V3 ray_dir, V3 n, float u, float v, V3 p,
                       V3 &attenuation)
{
    switch (m.kind) {
    case B200PT_MAT_LAMBERTIAN: {
        if (dot(ray_dir, n) < 0) {
            V3 value;
            float alpha;
            tex_eval(s, m.texture, u, v, p, value, alpha);
            attenuation = value * (1.0f / PT_PI_F);
        } else {
            attenuation = v3(0.f, 0.f, 0.f);
        }
        return true;
    }
    case B200PT_MAT_METAL: // a cosine lobe with constant albedo, no facing test (material.h:90-98) (sic)
        attenuation = v3(m.color[0], m.color[1], m.color[2]) * (1.0f / PT_PI_F);
        return true;
    case B200PT_MAT_DIELECTRIC:
        attenuation = v3(1.f, 1.f, 1.f);
        return true;
    case B200PT_MAT_ISOTROPIC: {
        V3 value;
        float alpha;
        tex_eval(s, m.texture, u, v, p, value, alpha);
        attenuation = value; // no 1/4pi (material.h:259) (sic)
        return true;
    }
    default:
        return false;
    }
}

// material::value: the pdf the material's own sampler has for `direction`
PT_HD float mat_value(int kind, V3 n, V3 direction)
{
    if (kind == B200PT_MAT_LAMBERTIAN || kind == B200PT_MAT_METAL) {
        // cosine_pdf::value (pdf.h:18-28); uvw.w() = unit_vector(n)
        float cosine = dot(unit_vector(direction), unit_vector(n));
        return cosine > 0 ? cosine * (1.0f / PT_PI_F) : 0.f;
    }
    if (kind == B200PT_MAT_ISOTROPIC)
        return (float)(1 / (4 * 3.14159265358979323846)); // random_pdf::value (pdf.h:41-44)
    return 0.f;                                          // void_pdf (pdf.h:71-74): dielectric, diffuse_light
}

PT_HD V3 random_in_unit_sphere(float r1, float r2, float r3)
{
    // random.h:17-24: a point INSIDE the ball (radius r3^(1/3)), y-up spherical coordinates
    float u = r1 * 2 * PT_PI_F;
    float v = acosf(2 * r2 - 1);
    float w = powf(r3, 1.0f / 3.0f);
    return v3(cosf(u) * sinf(v) * w, cosf(v) * w, sinf(u) * sinf(v) * w);
}

PT_HD V3 random_cosine_direction(float r1, float r2)
{
    // random.h:35-44
    float z = sqrtf(1 - r2);
    float phi = 2 * PT_PI_F * r1;
    float sq = sqrtf(r2);
    return v3(cosf(phi) * sq, sinf(phi) * sq, z);
}

PT_HD V3 reflect(V3 v, V3 n) { return v - 2 * dot(v, n) * n; }

PT_HD bool refract(V3 v, V3 n, float ni_over_nt, V3 &refracted)
{
    // helpers.h:24-36
    V3 uv = unit_vector(v);
    float dt = dot(uv, n);
    float discriminant = 1.0f - ni_over_nt * ni_over_nt * (1 - dt * dt);
    if (discriminant > 0) {
        refracted = ni_over_nt * (uv - n * dt) - n * sqrtf(discriminant);
        return true;
    }
    return false;
}

PT_HD float schlick(float cosine, float ref_idx)
{
    float r0 = (1 - ref_idx) / (1 + ref_idx);
    r0 = r0 * r0;
    float c = 1 - cosine;
    return r0 + (1 - r0) * (c * c * c * c * c);
}

// material::generate
PT_HD V3 mat_generate(const b200pt_material &m, V3 ray_dir, V3 n, float r1, float r2, float r3)
{
    if (m.kind == B200PT_MAT_LAMBERTIAN || m.kind == B200PT_MAT_METAL) {
        Onb uvw = onb_from_w(n);
        return onb_local(uvw, random_cosine_direction(r1, r2));
    }
    if (m.kind == B200PT_MAT_DIELECTRIC) {
        // material.h:125-167
        V3 outward_normal;
        V3 reflected = reflect(ray_dir, n);
        float ni_over_nt, cosine, reflect_prob;
        V3 refracted = v3(0.f, 0.f, 0.f);
        float dlen = length(ray_dir);
        if (dot(ray_dir, n) > 0) {
            outward_normal = -n;
            ni_over_nt = m.ior;
            cosine = m.ior * dot(ray_dir, n) / dlen;
        } else {
            outward_normal = n;
            ni_over_nt = 1.0f / m.ior;
            cosine = -dot(ray_dir, n) / dlen;
        }
        if (refract(ray_dir, outward_normal, ni_over_nt, refracted))
            reflect_prob = schlick(cosine, m.ior);
        else
            reflect_prob = 1.0f;
        return r1 < reflect_prob ? reflected : refracted;
    }
    // isotropic: random_pdf::generate; diffuse_light: void_pdf::generate (pdf.h:45-48,75-78)
    return random_in_unit_sphere(r1, r2, r3);
}

// ---- lights (hittable::random / pdf_value through the instance wrapper, primitive.h:319-342) ------------------

// sphere::random (primitive.h:51-58) + random_to_sphere (random.h:46-55)
PT_HD_COLD V3 sphere_random(V3 center, float radius, V3 lo, float r1, float r2)
{
    V3 direction = center - lo;
    float distance_squared = squared_length(direction);
    Onb uvw = onb_from_w(direction);
    float z = 1 + r2 * (sqrtf(1 - radius * radius / distance_squared) - 1);
    float phi = 2 * PT_PI_F * r1;
    float sq = sqrtf(1 - z * z);
    return onb_local(uvw, v3(cosf(phi) * sq, sinf(phi) * sq, z));
}

PT_HD V3 light_random(const InstanceData &I, const float *fwd, V3 o, float r1, float r2)
{
    V3 lo = xform_point(I.inv, o);
    V3 dl;
    int kind = meta_kind(I.meta);
    if (meta_medium(I.meta) || kind == B200PT_PRIM_BOX) {
        dl = v3(1.f, 0.f, 0.f); // hittable::random default (hittable.h:28): box and constant_medium do not override
    } else if (kind == B200PT_PRIM_RECT) {
        // rect::random (primitive.h:172-178)
        V3 pt = shuffle(v3(I.p[0] + r1 * (I.p[2] - I.p[0]), I.p[4], I.p[1] + r2 * (I.p[3] - I.p[1])), meta_plane(I.meta));
        dl = pt - lo;
    } else {
        dl = sphere_random(v3(I.p[0], I.p[1], I.p[2]), I.p[3], lo, r1, r2);
    }
    return xform_vector(fwd, dl);
}

// rec: the light's own hit record for the ray (o, v) over (0.001, FLT_MAX), valid when the pdf is > 0 -- the same
// record instance::hit would produce, so a shadow query along (o, v) need not intersect this light again.
PT_HD float light_pdf_value(const InstanceData &I, V3 o, V3 v, PrimHit &rec)
{
    int kind = meta_kind(I.meta);
    if (meta_medium(I.meta) || kind == B200PT_PRIM_BOX)
        return 0.f; // hittable::pdf_value default (hittable.h:27)
    V3 lo = xform_point(I.inv, o);
    V3 lv = xform_vector(I.inv, v);
    if (kind == B200PT_PRIM_RECT) {
        // rect::pdf_value (primitive.h:156-171): all lengths are LOCAL-space lengths (sic)
        if (hit_rect(meta_plane(I.meta), meta_flip(I.meta), I.p[0], I.p[1], I.p[2], I.p[3], I.p[4], lo, lv, 0.001f,
                     PT_FLT_MAX, rec)) {
            float area = (I.p[2] - I.p[0]) * (I.p[3] - I.p[1]);
            float vlen = length(lv);
            float d = rec.t * vlen;
            float distance_squared = d * d;
            float cosine = fabsf(dot(lv, rec.n) / vlen);
            return distance_squared / (cosine * area);
        }
        return 0.f;
    }
    // sphere::pdf_value (primitive.h:37-50)
    V3 center = v3(I.p[0], I.p[1], I.p[2]);
    float radius = I.p[3];
    if (hit_sphere(center, radius, lo, lv, 0.001f, PT_FLT_MAX, rec)) {
        float cos_theta_max = sqrtf(1 - radius * radius / squared_length(center - lo));
        float solid_angle = 2 * PT_PI_F * (1 - cos_theta_max);
        return 1 / solid_angle;
    }
    return 0.f;
}

PT_HD float light_pdf_value(const InstanceData &I, V3 o, V3 v)
{
    PrimHit rec;
    return light_pdf_value(I, o, v, rec);
}

PT_HD float power_heuristic(float f, float g)
{
    // helpers.h:138-144 with nf = ng = 1, pow = 2
    float f2 = f * f;
    return f2 / (f2 + g * g);
}

// ---- one next-event-estimation sample (integrator.h:221-265) --------------------------------------------------
// Split around the shadow ray so that a kernel can keep the ray in flight between the two halves:
//   nee_prepare  light pick, light sample, pdfs, MIS weight -> ray direction + everything but the emission
//   nee_finish   emission of whatever the shadow ray hit first -> the sample's contribution
// W = attenuation * beta at the vertex.  The contribution is before the division by light_samples.
struct NeeSample {
    V3 dir;  // shadow ray direction, un-normalised: the light point minus p (integrator.h:235)
    V3 coef; // attenuation * beta * weight_l / light_pdf_l * max(cos_l, 0)
    float pick_pdf;
    int known_inst; // >= 0: the sampled light is an emitter and `known` is its hit record along dir (light_pdf_value)
    PrimHit known;
};

PT_HD void nee_prepare(const PtSceneView &s, const RngKey &key, int bounce, int k, V3 p, V3 n, int mat_kind, V3 W,
                       NeeSample &out)
{
    F4 r = rng4(key, stage_word(bounce, RNG_NEE, k), 0);
    int li = (int)(r.x * s.n_lights); // World::get_random_light (world.h:31-35)
    if (li >= s.n_lights)
        li = s.n_lights - 1;
    const int light_inst = s.lights[li];
    out.pick_pdf = (float)s.n_lights; // (sic) integrator.h:224
    InstanceData L;
    load_instance(s.inst, light_inst, L);
    V3 dir = light_random(L, s.fwd[light_inst].m, p, r.y, r.z);
    float cos_l = dot(unit_vector(dir), unit_vector(n));
    float light_pdf_l = light_pdf_value(L, p, dir, out.known);
    out.known_inst = light_pdf_l > 0.f && meta_emitter(L.meta) ? light_inst : -1;
    float scatter_pdf_l = mat_value(mat_kind, n, dir);
    float weight_l = power_heuristic(light_pdf_l, scatter_pdf_l);
    float dropoff = fmaxf(cos_l, 0.0f);
    V3 c = W * weight_l;
    c = div_fast(c, light_pdf_l);
    c = c * dropoff;
    out.dir = dir;
    out.coef = c;
}

// Every light sample of a vertex is void (nee_is_void) when all lights lie below the vertex's horizon: the sampled
// direction points at a point of the light (rect::random, sphere::random), n . dir is then at most the largest
// n . (corner - p) over the corners of the light's world bounds, cos_l <= 0 and the coefficient is 0 (or NaN) for any
// material (`dropoff`, integrator.h:240).  The margin (1e-4 of the sum of the term magnitudes, against rounding of
// ~1e-6 in the computed cos_l) keeps the test conservative: a vertex it rejects has no sample that the full
// computation would keep, so the film does not change.  NaNs in p or n fail the comparison: no skip.
PT_HD bool lights_below_horizon(const PtSceneView &s, V3 p, V3 n)
{
    if (s.n_light_boxes == 0)
        return false;
    for (int l = 0; l < s.n_light_boxes; l++) {
        const PtBox &b = s.light_boxes[l];
        float m = 0.f, mag = 0.f;
        const float pp[3] = {p.x, p.y, p.z}, nn[3] = {n.x, n.y, n.z};
        for (int k = 0; k < 3; k++) {
            const float a = nn[k] * (b.lo[k] - pp[k]), c = nn[k] * (b.hi[k] - pp[k]);
            m += fmaxf(a, c);
            mag += fmaxf(fabsf(a), fabsf(c));
        }
        if (!(m < -1e-4f * mag))
            return false;
    }
    return true;
}

// a sample whose coefficient is zero or NaN contributes nothing whatever its shadow ray meets (nee_finish turns
// a NaN product into zero): no ray needs tracing.  The reference's ray counter still counts it (k_tally).
PT_HD bool nee_is_void(const NeeSample &ns)
{
    return is_nan(ns.coef) || (ns.coef.x == 0.f && ns.coef.y == 0.f && ns.coef.z == 0.f);
}

// lrec: the closest hit of the shadow ray; it starts at the un-offset hit point and t is in units of |dir|
// (integrator.h:235,246) (sic).  Whatever is hit first contributes its emission -- `true ||` (integrator.h:250) (sic)
PT_HD V3 nee_finish(const PtSceneView &s, const NeeSample &ns, bool hit, const Hit &lrec)
{
    if (!hit)
        return v3(0.f, 0.f, 0.f);
    const b200pt_material &lm = s.mats[meta_material(s.inst[lrec.inst].meta)];
    V3 light_emission = mat_emitted(s, lm, ns.dir, lrec.n, lrec.u, lrec.v, lrec.p);
    V3 c = ns.coef * light_emission;
    c = div_fast(c, ns.pick_pdf);
    if (is_nan(c))
        return v3(0.f, 0.f, 0.f);
    return c;
}

PT_HD MediumRng nee_medium_rng(const RngKey &key, int bounce, int k)
{
    MediumRng mr;
    mr.key = key;
    mr.stage = stage_word(bounce, RNG_NEE_MEDIUM, k);
    mr.xi = -1.f;
    return mr;
}

template <bool COUNT, bool MEDIA = true>
PT_HD V3 nee_sample(const PtSceneView &s, const RenderParams &rp, const RngKey &key, int bounce, int k, V3 p, V3 n,
                    int mat_kind, V3 W, TraceCounters *counters)
{
    NeeSample ns;
    nee_prepare(s, key, bounce, k, p, n, mat_kind, W, ns);
    if (nee_is_void(ns))
        return v3(0.f, 0.f, 0.f);
    MediumRng mr = nee_medium_rng(key, bounce, k);
    Hit lrec;
    bool hit = trace_shadow<COUNT, MEDIA>(s, p, ns.dir, 0.001f, PT_FLT_MAX, mr, lrec, counters, ns.known_inst, &ns.known);
    return nee_finish(s, ns, hit, lrec);
}

// ---- the vertex: emission, NEE request, scattering, Russian roulette, throughput (integrator.h:193-323) ------

struct PathVertexIn {
    V3 o, d;        // the ray that arrived
    V3 beta;        // throughput
    V3 attenuation; // value left by the previous vertex (integrator.h:181)
    float last_bsdf_pdf;
};

struct PathVertexOut {
    V3 add;         // radiance to add to the sample now
    bool want_nee;  // run light_samples NEE samples with W, at (p, n, mat_kind)
    V3 W;
    int mat_kind;
    bool cont;      // path continues with the fields below
    V3 o, d, beta, attenuation;
    float last_bsdf_pdf;
};

// QUEUE: the shade queue the vertex came from (1 lambertian|metal, 2 dielectric, 3 diffuse_light, 4 isotropic --
// ShadeKind in wavefront.cuh) or -1 when unknown (CPU restatement).  The kernels are instantiated per queue; telling
// the compiler which materials can occur removes the other materials' code from each instance.
template <int QUEUE> PT_HD void assume_queue(int mat_kind)
{
#if defined(__CUDA_ARCH__)
    if (QUEUE == 1)
        __builtin_assume(mat_kind == B200PT_MAT_LAMBERTIAN || mat_kind == B200PT_MAT_METAL);
    if (QUEUE == 2)
        __builtin_assume(mat_kind == B200PT_MAT_DIELECTRIC);
    if (QUEUE == 3)
        __builtin_assume(mat_kind == B200PT_MAT_DIFFUSE_LIGHT);
    if (QUEUE == 4)
        __builtin_assume(mat_kind == B200PT_MAT_ISOTROPIC);
#endif
}

template <int QUEUE = -1>
PT_HD void shade_hit_ineept(const PtSceneView &s, const RenderParams &rp, const PathVertexIn &in, const Hit &rec, F4 rnd,
                            PathVertexOut &out)
{
    const uint32_t meta = s.inst[rec.inst].meta;
    const b200pt_material m = s.mats[meta_material(meta)]; // a copy: the assumption below sticks to the value
    assume_queue<QUEUE>(m.kind);
    out.add = v3(0.f, 0.f, 0.f);
    out.cont = false;
    out.mat_kind = m.kind;
    V3 attenuation = in.attenuation;
    bool did_scatter = mat_scatter(s, m, in.d, rec.n, rec.u, rec.v, rec.p, attenuation);
    float cos_i = fabsf(dot(unit_vector(in.d), unit_vector(rec.n))); // cosine of the INCOMING ray (sic)
    V3 hit_emission = mat_emitted(s, m, in.d, rec.n, rec.u, rec.v, rec.p);
    if (squared_length(hit_emission) > 0.000001f) {
        if (in.last_bsdf_pdf <= 0) {
            out.add += in.beta * hit_emission;
        } else {
            // hittable_pdf(rec.primitive, r.origin()).value(rec.p): the hit POINT is passed as direction (sic, Q3)
            InstanceData I;
            load_instance(s.inst, rec.inst, I);
            float weight = power_heuristic(in.last_bsdf_pdf, light_pdf_value(I, in.o, rec.p));
            out.add += (in.beta * hit_emission) * weight;
        }
    }
    // NEE is skipped when it cannot contribute (integrator.h:248); the rays the reference would still cast are
    // accounted for by the caller
    out.want_nee = length(attenuation) > 0.0001f && s.n_lights > 0;
    out.W = attenuation * in.beta;

    if (did_scatter) {
        V3 dir = mat_generate(m, in.d, rec.n, rnd.x, rnd.y, rnd.z);
        float scatter_pdf_s = mat_value(m.kind, rec.n, dir);
        V3 beta = in.beta;
        float pmax = fmaxf(beta.x, fmaxf(beta.y, beta.z));
        if (rp.russian_roulette && pmax <= 1 && 0.001f < pmax) {
            if (rnd.w > pmax)
                return;
            beta = beta * (1 / pmax);
        }
        if (rp.only_direct)
            return;
        if (scatter_pdf_s < 0.0000001f)
            return;
        beta = beta * (attenuation * cos_i / scatter_pdf_s); // (sic, Q4)
        // Q12: a path whose throughput is exactly zero can add nothing any more; the reference keeps tracing it
        if (beta.x == 0.f && beta.y == 0.f && beta.z == 0.f)
            return;
        out.cont = true;
        out.o = rec.p + rp.normal_offset * rec.n;
        out.d = dir;
        out.beta = beta;
        out.attenuation = attenuation;
        out.last_bsdf_pdf = scatter_pdf_s;
    } else {
        out.add += in.beta * hit_emission; // emission counted a second time (sic, Q2)
    }
}

// RecursivePT::color unrolled: radiance = sum over vertices of (product of attenuations) * emitted
// (integrator.h:29-72).  depth = index of this vertex (0 = first hit).
template <int QUEUE = -1>
PT_HD void shade_hit_rpt(const PtSceneView &s, const RenderParams &rp, const PathVertexIn &in, const Hit &rec, int depth,
                         F4 rnd, PathVertexOut &out)
{
    const uint32_t meta = s.inst[rec.inst].meta;
    const b200pt_material m = s.mats[meta_material(meta)];
    assume_queue<QUEUE>(m.kind);
    out.want_nee = false;
    out.cont = false;
    out.mat_kind = m.kind;
    out.W = v3(0.f, 0.f, 0.f);
    V3 emitted = mat_emitted(s, m, in.d, rec.n, rec.u, rec.v, rec.p);
    out.add = in.beta * emitted;
    V3 attenuation = in.attenuation;
    if (depth < rp.max_bounces && mat_scatter(s, m, in.d, rec.n, rec.u, rec.v, rec.p, attenuation)) {
        V3 beta = in.beta * attenuation;
        if (beta.x == 0.f && beta.y == 0.f && beta.z == 0.f)
            return;
        out.cont = true;
        out.o = rec.p; // no normal offset in this integrator (integrator.h:44)
        out.d = mat_generate(m, in.d, rec.n, rnd.x, rnd.y, rnd.z);
        out.beta = beta;
        out.attenuation = attenuation;
        out.last_bsdf_pdf = -1.f;
    }
}

// miss: background lookup (integrator.h:63-69,325-335).  u,v only matter for image textures; p = unit direction.
PT_HD V3 shade_miss(const PtSceneView &s, V3 d, V3 beta)
{
    V3 value;
    float alpha;
    const V3 ud = unit_vector(d);
    float u = 0.f, v = 0.f;
    if (s.n_images > 0) {
        // integrator.h:67-68,329-330 with `#define TAU 2 * M_PI` (random.h:7): x / TAU is (x / 2) * M_PI (sic, Q10)
        u = (float)((3.14159265358979323846 + (double)atan2f(ud.y, ud.x)) / 2 * 3.14159265358979323846);
        v = (float)((double)acosf(ud.z) / 3.14159265358979323846);
    }
    tex_eval(s, s.background, u, v, ud, value, alpha);
    return beta * value;
}

// camera::get_ray (camera.h:38-47) for film coordinates (s, t); rd = lens sample in the unit disk
PT_HD void camera_ray(const PtCamera &c, float sx, float ty, float r1, float r2, V3 &o, V3 &d)
{
    V3 offset = v3(0.f, 0.f, 0.f);
    if (c.lens_radius > 0.f) {
        // random_in_unit_disk (random.h:26-33)
        float u = r1 * 2 * PT_PI_F;
        float v = sqrtf(r2);
        V3 rd = c.lens_radius * v3(cosf(u) * v, sinf(u) * v, 0.f);
        offset = c.u * rd.x + c.v * rd.y;
    }
    o = c.origin + offset;
    d = c.lower_left + sx * c.horizontal + ty * c.vertical - c.origin - offset;
}

} // namespace pt
