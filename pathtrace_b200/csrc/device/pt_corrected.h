// pt_corrected.h -- the opt-in "physically corrected" vertex logic (config "physically_corrected": true).
//
// NOT the reference's estimator and never used for parity: the code at HEAD counts emission twice, weights the
// throughput with the INCOMING cosine (infinite variance), divides NEE by n_lights instead of multiplying, lets
// `metal` be a diffuse lobe and ends every path at a dielectric.  This header is what those lines
// (material.h:76-175, integrator.h:201-218,254,305) are meant to compute -- a standard unidirectional path tracer:
//
//   lambertian   f = albedo / pi, cosine-sampled about the face-forwarded normal
//   metal        mirror reflection perturbed by fuzz * (point in the unit ball) (material.h:76-89), tint = albedo;
//                specular: no NEE, the next emitter hit counts in full
//   dielectric   Schlick reflect / refract (material.h:125-167), specular, the path goes on
//   isotropic    f = albedo / 4pi, uniform direction (constant_medium's phase function)
//   emission     counted ONCE, with the power-heuristic weight against the light-sampling pdf when the previous vertex
//                could have sampled it (NEE on, non-specular), in full otherwise
//   NEE          one of the n lights uniformly (probability 1/n), a point on it (rect: uniform area, sphere: cone),
//                contribution beta f cos_s Le / (p_light / n) * w_mis, only if the SAMPLED light is what the shadow ray
//                meets first; pdfs are solid-angle measures evaluated as the reference evaluates them (instance-local
//                lengths: exact for the rigid transforms every light of the shipped scenes has)
//   throughput   beta *= f cos_s / p_s  (= albedo for the two diffuse lobes, tint for metal, 1 for glass)
//
// Shared by the kernels and the CPU restatement like the rest of device/, so the GPU is checked against the
// restatement sample for sample; its correctness as an ESTIMATOR is checked by comparing NEE + MIS with plain path
// tracing (both unbiased for the same integral) in tests/test_gpu_corrected.py.
#pragma once
#include "pt_shade.h"

namespace pt {

// nee_w.w of the corrected mode: material kind | (the shading normal is the flipped geometric normal) << 8
PT_HD int corrected_kind_word(int mat_kind, bool flipped) { return mat_kind | (flipped ? 0x100 : 0); }

// uniform direction on the unit sphere (the normalised counterpart of random.h:17-24)
PT_HD V3 uniform_sphere_direction(float r1, float r2)
{
    float z = 1 - 2 * r2;
    float s = sqrtf(fmaxf(0.f, 1 - z * z));
    float phi = 2 * PT_PI_F * r1;
    return v3(cosf(phi) * s, z, sinf(phi) * s);
}

template <int QUEUE = -1>
PT_HD void shade_hit_corrected(const PtSceneView &s, const RenderParams &rp, const PathVertexIn &in, const Hit &rec,
                               F4 rnd, PathVertexOut &out)
{
    const bool nee = rp.integrator == B200PT_INTEGRATOR_INEEPT;
    const uint32_t meta = s.inst[rec.inst].meta;
    const b200pt_material m = s.mats[meta_material(meta)];
    assume_queue<QUEUE>(m.kind);
    out.add = v3(0.f, 0.f, 0.f);
    out.cont = false;
    out.want_nee = false;
    out.W = v3(0.f, 0.f, 0.f);
    const bool flipped = !(dot(in.d, rec.n) < 0);
    const V3 n_ff = flipped ? -rec.n : rec.n;
    out.mat_kind = corrected_kind_word(m.kind, flipped);

    // emission, once
    V3 Le = mat_emitted(s, m, in.d, rec.n, rec.u, rec.v, rec.p);
    if (Le.x != 0.f || Le.y != 0.f || Le.z != 0.f) {
        float w = 1.f;
        if (nee && in.last_bsdf_pdf > 0 && meta_light(meta) && s.n_lights > 0) {
            InstanceData I;
            load_instance(s.inst, rec.inst, I);
            float p_l = light_pdf_value(I, in.o, in.d) / (float)s.n_lights;
            w = power_heuristic(in.last_bsdf_pdf, p_l);
        }
        out.add = (in.beta * Le) * w;
    }
    if (m.kind == B200PT_MAT_DIFFUSE_LIGHT)
        return;

    V3 dir, tint;
    float pdf; // < 0: specular
    V3 value;
    float alpha;
    if (m.kind == B200PT_MAT_LAMBERTIAN) {
        tex_eval(s, m.texture, rec.u, rec.v, rec.p, value, alpha);
        tint = value;
        Onb uvw = onb_from_w(n_ff);
        dir = onb_local(uvw, random_cosine_direction(rnd.x, rnd.y));
        float c = dot(unit_vector(dir), unit_vector(n_ff));
        pdf = c > 0 ? c * (1.0f / PT_PI_F) : 0.f;
        if (!(pdf > 0))
            return;
    } else if (m.kind == B200PT_MAT_METAL) {
        tint = v3(m.color[0], m.color[1], m.color[2]);
        dir = reflect(unit_vector(in.d), n_ff) + m.power * random_in_unit_sphere(rnd.x, rnd.y, rnd.z); // power = fuzz
        if (!(dot(dir, n_ff) > 0))
            return; // scattered below the surface: absorbed (material.h:76-89 as written for book 1)
        pdf = -1.f;
    } else if (m.kind == B200PT_MAT_DIELECTRIC) {
        tint = v3(1.f, 1.f, 1.f);
        dir = mat_generate(m, in.d, rec.n, rnd.x, rnd.y, rnd.z);
        pdf = -1.f;
    } else { // isotropic
        tex_eval(s, m.texture, rec.u, rec.v, rec.p, value, alpha);
        tint = value;
        dir = uniform_sphere_direction(rnd.x, rnd.y);
        pdf = (float)(1 / (4 * 3.14159265358979323846));
    }
    if (pdf > 0 && nee && s.n_lights > 0) {
        out.want_nee = true;
        out.W = in.beta * tint;
    }
    V3 beta = in.beta * tint;
    float pmax = fmaxf(beta.x, fmaxf(beta.y, beta.z));
    if (!(pmax > 0))
        return;
    if (rp.russian_roulette && pmax < 1) {
        if (rnd.w > pmax)
            return;
        beta = beta * (1 / pmax);
    }
    if (rp.only_direct)
        return;
    out.cont = true;
    // leave on the side the new direction points to (a refracted ray starts below the surface)
    const bool medium = m.kind == B200PT_MAT_ISOTROPIC;
    out.o = medium ? rec.p : rec.p + (dot(dir, n_ff) >= 0 ? rp.normal_offset : -rp.normal_offset) * n_ff;
    out.d = dir;
    out.beta = beta;
    out.attenuation = tint;
    out.last_bsdf_pdf = pdf;
}

// One light sample of the corrected estimator.  kind_word: corrected_kind_word(); W = beta * albedo at the vertex.
PT_HD void nee_prepare_corrected(const PtSceneView &s, const RngKey &key, int bounce, int k, V3 p, V3 n, int kind_word,
                                 V3 W, NeeSample &out)
{
    F4 r = rng4(key, stage_word(bounce, RNG_NEE, k), 0);
    int li = (int)(r.x * s.n_lights);
    if (li >= s.n_lights)
        li = s.n_lights - 1;
    const int light_inst = s.lights[li];
    const int mat_kind = kind_word & 0xff;
    const V3 n_ff = (kind_word & 0x100) ? -n : n;
    InstanceData L;
    load_instance(s.inst, light_inst, L);
    V3 dir = light_random(L, s.fwd[light_inst].m, p, r.y, r.z);
    float p_l = light_pdf_value(L, p, dir, out.known);
    out.known_inst = p_l > 0.f && meta_emitter(L.meta) ? light_inst : -1;
    out.pick_pdf = 1.f;
    out.dir = dir;
    out.coef = v3(0.f, 0.f, 0.f);
    if (out.known_inst < 0)
        return;
    float cos_s, p_s;
    V3 f;
    if (mat_kind == B200PT_MAT_ISOTROPIC) {
        cos_s = 1.f;
        p_s = (float)(1 / (4 * 3.14159265358979323846));
        f = W * p_s;
    } else {
        cos_s = fmaxf(dot(unit_vector(dir), unit_vector(n_ff)), 0.f);
        p_s = cos_s * (1.0f / PT_PI_F);
        f = W * (1.0f / PT_PI_F);
    }
    const float p_pick = p_l / (float)s.n_lights;
    const float w = power_heuristic(p_pick, p_s);
    out.coef = f * (cos_s * w / p_pick);
}

// lrec: the emitter the shadow ray met first (trace_shadow).  Only the SAMPLED light may answer for its own pdf.
PT_HD V3 nee_finish_corrected(const PtSceneView &s, const NeeSample &ns, bool hit, const Hit &lrec)
{
    if (!hit || lrec.inst != ns.known_inst)
        return v3(0.f, 0.f, 0.f);
    const b200pt_material &lm = s.mats[meta_material(s.inst[lrec.inst].meta)];
    V3 c = ns.coef * mat_emitted(s, lm, ns.dir, lrec.n, lrec.u, lrec.v, lrec.p);
    if (is_nan(c))
        return v3(0.f, 0.f, 0.f);
    return c;
}

template <bool COUNT, bool MEDIA = true>
PT_HD V3 nee_sample_corrected(const PtSceneView &s, const RngKey &key, int bounce, int k, V3 p, V3 n, int kind_word, V3 W,
                              TraceCounters *counters)
{
    NeeSample ns;
    nee_prepare_corrected(s, key, bounce, k, p, n, kind_word, W, ns);
    if (nee_is_void(ns))
        return v3(0.f, 0.f, 0.f);
    MediumRng mr = nee_medium_rng(key, bounce, k);
    Hit lrec;
    bool hit = trace_shadow<COUNT, MEDIA>(s, p, ns.dir, 0.001f, PT_FLT_MAX, mr, lrec, counters, ns.known_inst, &ns.known);
    return nee_finish_corrected(s, ns, hit, lrec);
}

} // namespace pt
