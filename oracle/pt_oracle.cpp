// oracle/pt_oracle.cpp -- TEST INFRASTRUCTURE ONLY (never linked into the product; only tests/, smoke() and
// bench.py's cpu_baseline leg may load it).
//
// CPU restatement of the hot path as ONE scalar loop per camera sample, in the shape of the reference's own
// control flow:
//     Tiled::compute        renderer.h:626-691   pixel/sample loop, framebuffer[j][i] += de_nan(col)
//     NEEIterative::color   integrator.h:176-339 bounce loop: hit -> scatter/emitted -> NEE x light_samples ->
//                                                generate -> Russian roulette -> throughput
//     RecursivePT::color    integrator.h:29-72
//     World::hit            world.h:17-20        (flat BVH here)
// It is built from the same `__host__ __device__` math headers as the CUDA kernels (pathtrace_b200/csrc/device/
// pt_*.h, compiled here by g++ -ffp-contract=off) and draws from the same Philox stream keyed by
// (pixel, sample, bounce, stage), so a GPU render and this loop agree per SAMPLE, not just statistically: it
// checks the wavefront plumbing (queues, compaction, sorting, batching, accumulation).  It is NOT the
// independent oracle: that is oracle/_ref/ref_driver, the unmodified reference, against which this restatement
// itself is pinned statistically on the CPU (tests/test_oracle_vs_reference.py) and the hit routine ray by ray.
// PARITY PIN: hit routine -- exact comparison with the reference on fixed ray batches; render -- statistical
// (clamped per-pixel z-scores) against the reference's integrator->color; no upstream golden vectors exist.
#include "../include/b200pt.h"
#include "../pathtrace_b200/csrc/device/pt_corrected.h"
#include "../pathtrace_b200/csrc/host/flat_scene.h"
#include "../pathtrace_b200/csrc/host/scene_pack.h"

#include <atomic>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

using namespace b200pt;
using namespace pt;

namespace {

thread_local std::string g_err;

struct OracleScene {
    FlatScene flat;
    PackedScene packed;
    PtSceneView view;
};

struct SampleCounts {
    long ext = 0, shadow = 0, reference = 0;
};

RenderParams make_params(const b200pt_config &c)
{
    RenderParams rp{};
    rp.width = c.width;
    rp.height = c.height;
    rp.integrator = c.integrator;
    rp.max_bounces = c.max_bounces;
    rp.light_samples = c.integrator == B200PT_INTEGRATOR_INEEPT ? c.light_samples : 0;
    rp.light_group = 0; // the lane-group size of k_nee; the scalar loop below reproduces its summation order itself
    rp.russian_roulette = c.russian_roulette;
    rp.only_direct = c.only_direct_illumination;
    rp.corrected = c.physically_corrected ? 1 : 0;
    rp.normal_offset = c.normal_offset;
    rp.seed_lo = (uint32_t)(c.seed & 0xffffffffull);
    rp.seed_hi = (uint32_t)(c.seed >> 32);
    return rp;
}

// one camera sample = one call of integrator->color (renderer.h:648-670)
V3 sample_color(const PtSceneView &s, const RenderParams &rp, uint32_t pixel, long long sample, SampleCounts &cnt)
{
    RngKey key;
    key.seed_lo = rp.seed_lo;
    key.seed_hi = rp.seed_hi;
    key.pixel = pixel;
    key_set_sample(key, sample);
    F4 r = rng4(key, stage_word(0, RNG_CAMERA, 0), 0);
    int j = pixel / rp.width, i = pixel - j * rp.width;
    float u = (float)(i + (double)r.x) / (float)rp.width;
    float v = (float)(j + (double)r.y) / (float)rp.height;
    PathVertexIn in;
    camera_ray(s.cam, u, v, r.z, r.w, in.o, in.d);
    in.beta = v3(1.f, 1.f, 1.f);
    in.attenuation = v3(0.f, 0.f, 0.f);
    in.last_bsdf_pdf = -1.f;
    V3 sum = v3(0.f, 0.f, 0.f);
    const bool nee = rp.integrator == B200PT_INTEGRATOR_INEEPT;
    const int n_bounces = nee ? rp.max_bounces : rp.max_bounces + 1;
    for (int b = 0; b < n_bounces; b++) {
        MediumRng mr;
        mr.key = key;
        mr.stage = stage_word(b, RNG_EXT_MEDIUM, 0);
        mr.xi = -1.f;
        Hit rec;
        cnt.ext++;
        if (nee || b > 0)
            cnt.reference++;
        if (!trace_closest<false>(s, in.o, in.d, 0.001f, PT_FLT_MAX, mr, rec, nullptr)) {
            sum += shade_miss(s, in.d, in.beta);
            break;
        }
        F4 rnd = rng4(key, stage_word(b, RNG_SHADE, 0), 0);
        PathVertexOut out;
        // `attenuation` only survives from the previous vertex when a light is hit (material.h:187-191)
        PathVertexIn vin = in;
        if (s.mats[meta_material(s.inst[rec.inst].meta)].kind != B200PT_MAT_DIFFUSE_LIGHT)
            vin.attenuation = v3(0.f, 0.f, 0.f);
        if (rp.corrected)
            shade_hit_corrected(s, rp, vin, rec, rnd, out);
        else if (nee)
            shade_hit_ineept(s, rp, vin, rec, rnd, out);
        else
            shade_hit_rpt(s, rp, vin, rec, b, rnd, out);
        sum += out.add;
        if (nee)
            cnt.reference += rp.light_samples;
        if (out.want_nee) {
            // summed in the order of k_nee's butterfly over a lane group of P = light_samples rounded up to a power
            // of two (pairs, pairs of pairs, ...), so that the two sums can be compared bit for bit
            V3 v[32];
            V3 light = v3(0.f, 0.f, 0.f);
            int P = 1;
            while (P < rp.light_samples && P < 32)
                P *= 2;
            const bool tree = rp.light_samples <= 32;
            for (int k = 0; k < (tree ? P : rp.light_samples); k++) {
                V3 c = v3(0.f, 0.f, 0.f);
                if (k < rp.light_samples) {
                    c = rp.corrected ? nee_sample_corrected<false>(s, key, b, k, rec.p, rec.n, out.mat_kind, out.W, nullptr)
                                     : nee_sample<false>(s, rp, key, b, k, rec.p, rec.n, out.mat_kind, out.W, nullptr);
                    cnt.shadow++;
                }
                if (tree)
                    v[k] = c;
                else
                    light += c;
            }
            if (tree) {
                for (int off = 1; off < P; off <<= 1)
                    for (int i = 0; i < P; i += 2 * off)
                        v[i] = v[i] + v[i + off];
                light = v[0];
            }
            // k_nee adds nothing when the group sum is exactly zero; x + 0 = x leaves the bits alone either way
            sum += light / (float)rp.light_samples;
        }
        if (!out.cont)
            break;
        in.o = out.o;
        in.d = out.d;
        in.beta = out.beta;
        in.attenuation = out.attenuation;
        in.last_bsdf_pdf = out.last_bsdf_pdf;
    }
    return sum;
}

} // namespace

extern "C" {

const char *pt_oracle_last_error(void) { return g_err.c_str(); }

// the bare generator of pt_rng.h (known-answer tests): n x {counter[4], key[2]} -> n x 4 words
void pt_oracle_philox(const uint32_t *in, size_t n, uint32_t *out)
{
    for (size_t i = 0; i < n; i++) {
        U4 c;
        c.x = in[6 * i], c.y = in[6 * i + 1], c.z = in[6 * i + 2], c.w = in[6 * i + 3];
        const U4 r = philox4x32_10(c, in[6 * i + 4], in[6 * i + 5]);
        out[4 * i] = r.x, out[4 * i + 1] = r.y, out[4 * i + 2] = r.z, out[4 * i + 3] = r.w;
    }
}

// the four uniforms rng4() hands out for (seed; pixel, global sample, stage word, index) and the stage word itself
void pt_oracle_rng4(uint64_t seed, uint32_t pixel, int64_t sample, int bounce, int stage, int k, uint32_t idx, float *out4,
                    uint32_t *stage_word_out)
{
    RngKey key;
    key.seed_lo = (uint32_t)(seed & 0xffffffffull);
    key.seed_hi = (uint32_t)(seed >> 32);
    key.pixel = pixel;
    key_set_sample(key, sample);
    const uint32_t sw = stage_word(bounce, stage, k);
    const F4 r = rng4(key, sw, idx);
    out4[0] = r.x, out4[1] = r.y, out4[2] = r.z, out4[3] = r.w;
    if (stage_word_out)
        *stage_word_out = sw;
}

// asset_dir: where png textures are looked up after the working directory (may be null)
void *pt_oracle_scene_create2(const char *scene_json, int film_w, int film_h, const char *asset_dir)
{
    try {
        OracleScene *sc = new OracleScene;
        sc->flat = compile_scene(Json::parse(scene_json), film_w, film_h, asset_dir ? asset_dir : "");
        sc->packed = pack_scene(sc->flat);
        sc->view = sc->packed.host_view(sc->flat);
        return sc;
    } catch (const std::exception &e) {
        g_err = e.what();
        return nullptr;
    }
}

void *pt_oracle_scene_create(const char *scene_json, int film_w, int film_h)
{
    return pt_oracle_scene_create2(scene_json, film_w, film_h, nullptr);
}

void pt_oracle_scene_destroy(void *scene) { delete (OracleScene *)scene; }

// World::hit for n rays; same contract as b200pt_trace_batch.  shadow != 0: the NEE shadow query instead
// (trace_shadow: a record only when the first thing met is an emitter), and emitter_flags[i] (optional, n_inst
// bytes) reports which instances are emitters -- tests check it against World::hit filtered by those flags.
static int trace_impl(void *scene, const float *rays, size_t n, const float *medium_xi, b200pt_hit *hits, int shadow)
{
    OracleScene *sc = (OracleScene *)scene;
    PtSceneView view = sc->view;
    view.emit_needs_n = 1;
    view.emit_needs_p = 1; // report p and n of shadow hits too (renders skip it when every emitter has a constant texture)
    for (size_t i = 0; i < n; i++) {
        const float *q = rays + i * 8;
        MediumRng mr;
        mr.key.seed_lo = mr.key.seed_hi = 0;
        mr.key.pixel = (uint32_t)(i & 0xffffffffull);
        mr.key.sample = (uint32_t)(i >> 32);
        mr.key.sample_hi = 0;
        mr.stage = stage_word(0, RNG_BATCH_MEDIUM, 0);
        mr.xi = medium_xi ? medium_xi[i] : -1.f;
        Hit h;
        bool ok;
        if (shadow == 2) {
            // no tree at all: every instance in index order (the closest hit and its tie rule do not depend on the
            // order of the tests) -- the yardstick for the tree the product builds and walks
            Traversal T;
            TraversalStack T_stack;
            T.stk = &T_stack;
            T.init(v3(q[0], q[1], q[2]), v3(q[3], q[4], q[5]), q[6], q[7]);
            for (int idx = 0; idx < view.n_inst; idx++)
                T.leaf_test<false, true>(view, idx, mr, nullptr);
            ok = T.result<true>(view, h);
        } else {
            ok = shadow ? trace_shadow<false>(view, v3(q[0], q[1], q[2]), v3(q[3], q[4], q[5]), q[6], q[7], mr, h, nullptr)
                        : trace_closest<false>(view, v3(q[0], q[1], q[2]), v3(q[3], q[4], q[5]), q[6], q[7], mr, h, nullptr);
        }
        b200pt_hit &o = hits[i];
        memset(&o, 0, sizeof(o));
        o.instance = -1;
        if (ok) {
            o.instance = h.inst;
            o.t = h.t;
            o.p[0] = h.p.x, o.p[1] = h.p.y, o.p[2] = h.p.z;
            o.n[0] = h.n.x, o.n[1] = h.n.y, o.n[2] = h.n.z;
            o.u = h.u;
            o.v = h.v;
        }
    }
    return 0;
}

int pt_oracle_trace(void *scene, const float *rays, size_t n, const float *medium_xi, b200pt_hit *hits)
{
    return trace_impl(scene, rays, n, medium_xi, hits, 0);
}

int pt_oracle_trace_shadow(void *scene, const float *rays, size_t n, const float *medium_xi, b200pt_hit *hits)
{
    return trace_impl(scene, rays, n, medium_xi, hits, 1);
}

int pt_oracle_trace_linear(void *scene, const float *rays, size_t n, const float *medium_xi, b200pt_hit *hits)
{
    return trace_impl(scene, rays, n, medium_xi, hits, 2);
}

// The horizon test k_shade uses to leave a vertex out of the NEE queue (pt_shade.h lights_below_horizon) against
// the full computation: for n vertices (pn: n x 6 = p, normal) draw `samples` light samples each with nee_prepare.
// out[0] = vertices the test rejects, out[1] = samples of rejected vertices that are NOT void (must be 0),
// out[2] = vertices the test keeps although all their samples are void (what the test misses; informational).
int pt_oracle_horizon_check(void *scene, const float *pn, size_t n, int samples, uint64_t *out)
{
    OracleScene *sc = (OracleScene *)scene;
    const PtSceneView &view = sc->view;
    out[0] = out[1] = out[2] = 0;
    for (size_t i = 0; i < n; i++) {
        const V3 p = v3(pn[6 * i], pn[6 * i + 1], pn[6 * i + 2]), nrm = v3(pn[6 * i + 3], pn[6 * i + 4], pn[6 * i + 5]);
        const bool rejected = lights_below_horizon(view, p, nrm);
        RngKey key;
        key.seed_lo = 17, key.seed_hi = 0;
        key.pixel = (uint32_t)i;
        key.sample = 0;
        key.sample_hi = 0;
        int live = 0;
        for (int k = 0; k < samples; k++) {
            NeeSample ns;
            nee_prepare(view, key, 0, k, p, nrm, B200PT_MAT_LAMBERTIAN, v3(1.f, 1.f, 1.f), ns);
            if (!nee_is_void(ns))
                live++;
        }
        if (rejected) {
            out[0]++;
            out[1] += (uint64_t)live;
        } else if (live == 0) {
            out[2]++;
        }
    }
    return 0;
}

// emitter flag of every instance (n_inst bytes); returns n_inst
int pt_oracle_emitters(void *scene, uint8_t *flags)
{
    OracleScene *sc = (OracleScene *)scene;
    for (int i = 0; i < sc->view.n_inst; i++)
        if (flags)
            flags[i] = meta_emitter(sc->view.inst[i].meta) ? 1 : 0;
    return sc->view.n_inst;
}

// The render loop.  rgb_sums: H x W x 3 float32 sums, row 0 = image bottom (Renderer::framebuffer).
// clamped_sum / clamped_sumsq: optional parity statistics (config.stats_clamp).  counts[3]: extension rays,
// shadow rays, reference-accounting rays.  pixel_begin/pixel_end restrict the work to a pixel range (timing
// samples); threads = worker count (rows are interleaved; results do not depend on it).
int pt_oracle_render(void *scene, const b200pt_config *config, float *rgb_sums, float *clamped_sum, float *clamped_sumsq,
                     int64_t counts[3], int threads, int64_t pixel_begin, int64_t pixel_end)
{
    OracleScene *sc = (OracleScene *)scene;
    const b200pt_config &c = *config;
    RenderParams rp = make_params(c);
    const int64_t npix = (int64_t)c.width * c.height;
    if (pixel_end <= 0 || pixel_end > npix)
        pixel_end = npix;
    if (pixel_begin < 0)
        pixel_begin = 0;
    if (threads < 1)
        threads = 1;
    std::atomic<long> t_ext{0}, t_shadow{0}, t_ref{0};
    std::atomic<int64_t> next{pixel_begin};
    const int64_t chunk = 256;
    auto worker = [&]() {
        SampleCounts cnt;
        while (true) {
            int64_t p0 = next.fetch_add(chunk);
            if (p0 >= pixel_end)
                break;
            int64_t p1 = p0 + chunk < pixel_end ? p0 + chunk : pixel_end;
            for (int64_t pix = p0; pix < p1; pix++) {
                float acc[3] = {0, 0, 0}, cs[3] = {0, 0, 0}, cq[3] = {0, 0, 0};
                for (int sidx = 0; sidx < c.samples; sidx++) {
                    V3 col = sample_color(sc->view, rp, (uint32_t)pix, (long long)c.sample_offset + sidx, cnt);
                    float x[3] = {is_nan(col.x) ? 0.f : col.x, is_nan(col.y) ? 0.f : col.y, is_nan(col.z) ? 0.f : col.z};
                    for (int k = 0; k < 3; k++) {
                        acc[k] += x[k];
                        if (c.stats_clamp > 0.f) {
                            float cx = x[k] < c.stats_clamp ? x[k] : c.stats_clamp; // two-sided: a Perlin albedo can be negative
                            cx = cx > -c.stats_clamp ? cx : -c.stats_clamp;
                            cs[k] += cx;
                            cq[k] += cx * cx;
                        }
                    }
                }
                for (int k = 0; k < 3; k++) {
                    rgb_sums[3 * pix + k] = acc[k];
                    if (clamped_sum)
                        clamped_sum[3 * pix + k] = cs[k];
                    if (clamped_sumsq)
                        clamped_sumsq[3 * pix + k] = cq[k];
                }
            }
        }
        t_ext += cnt.ext;
        t_shadow += cnt.shadow;
        t_ref += cnt.reference;
    };
    std::vector<std::thread> pool;
    for (int t = 1; t < threads; t++)
        pool.emplace_back(worker);
    worker();
    for (auto &t : pool)
        t.join();
    if (counts) {
        counts[0] = t_ext;
        counts[1] = t_shadow;
        counts[2] = t_ref;
    }
    return 0;
}

} // extern "C"
