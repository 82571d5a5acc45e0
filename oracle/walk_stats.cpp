// oracle/walk_stats.cpp -- TEST INFRASTRUCTURE / development tool, never linked into the product.
//
// Walk statistics of the CPU restatement (the same __host__ __device__ headers the kernels compile): slab tests and
// primitive tests per extension ray and per traced shadow ray on a small film of a scene, with the film's checksum.
// A change of the traversal (child order, culling, node layout) must leave the checksum alone -- the closest hit does
// not depend on the order of the tests -- and shows here as boxes / prims per ray before any GPU time is spent.
//
//   walk_stats <scene.json> [width height spp]
#include "../pathtrace_b200/csrc/device/pt_corrected.h"
#include "../pathtrace_b200/csrc/host/flat_scene.h"
#include "../pathtrace_b200/csrc/host/scene_pack.h"

#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <sstream>
#include <string>

using namespace b200pt;
using namespace pt;

int main(int argc, char **argv)
{
    if (argc < 2) {
        fprintf(stderr, "usage: walk_stats <scene.json> [width height spp]\n");
        return 2;
    }
    const int W = argc > 2 ? atoi(argv[2]) : 96, H = argc > 3 ? atoi(argv[3]) : 54, S = argc > 4 ? atoi(argv[4]) : 4;
    std::ifstream f(argv[1]);
    std::stringstream ss;
    ss << f.rdbuf();
    std::string dir = argv[1];
    dir = dir.substr(0, dir.find_last_of('/') == std::string::npos ? 0 : dir.find_last_of('/'));
    FlatScene flat = compile_scene(Json::parse(ss.str()), W, H, dir);
    PackedScene packed = pack_scene(flat);
    PtSceneView s = packed.host_view(flat);

    RenderParams rp{};
    rp.width = W, rp.height = H;
    rp.integrator = B200PT_INTEGRATOR_INEEPT;
    rp.max_bounces = 10, rp.light_samples = 4, rp.russian_roulette = 1, rp.normal_offset = 0.f;
    rp.seed_lo = 1234u;

    TraceCounters ext{0, 0}, sh{0, 0};
    unsigned long long n_ext = 0, n_sh = 0, n_sh_traced = 0;
    double checksum = 0;
    unsigned long long kinds[10][8] = {};
    for (int pix = 0; pix < W * H; pix++)
        for (int sidx = 0; sidx < S; sidx++) {
            RngKey key;
            key.seed_lo = rp.seed_lo, key.seed_hi = rp.seed_hi, key.pixel = (uint32_t)pix;
            key_set_sample(key, sidx);
            F4 r = rng4(key, stage_word(0, RNG_CAMERA, 0), 0);
            int j = pix / W, i = pix - j * W;
            PathVertexIn in;
            camera_ray(s.cam, (float)(i + (double)r.x) / (float)W, (float)(j + (double)r.y) / (float)H, r.z, r.w, in.o, in.d);
            in.beta = v3(1.f, 1.f, 1.f), in.attenuation = v3(0.f, 0.f, 0.f), in.last_bsdf_pdf = -1.f;
            V3 sum = v3(0.f, 0.f, 0.f);
            for (int b = 0; b < rp.max_bounces; b++) {
                MediumRng mr;
                mr.key = key, mr.stage = stage_word(b, RNG_EXT_MEDIUM, 0), mr.xi = -1.f;
                Hit rec;
                n_ext++;
                TraceCounters tc{0, 0};
                const bool hit = trace_closest<true>(s, in.o, in.d, 0.001f, PT_FLT_MAX, mr, rec, &tc);
                ext.boxes += tc.boxes, ext.prims += tc.prims;
                kinds[b][hit ? 1 + s.mats[meta_material(s.inst[rec.inst].meta)].kind : 0]++;
                if (!hit) {
                    sum += shade_miss(s, in.d, in.beta);
                    break;
                }
                F4 rnd = rng4(key, stage_word(b, RNG_SHADE, 0), 0);
                PathVertexOut out;
                PathVertexIn vin = in;
                if (s.mats[meta_material(s.inst[rec.inst].meta)].kind != B200PT_MAT_DIFFUSE_LIGHT)
                    vin.attenuation = v3(0.f, 0.f, 0.f);
                shade_hit_ineept(s, rp, vin, rec, rnd, out);
                sum += out.add;
                if (out.want_nee)
                    for (int k = 0; k < rp.light_samples; k++) {
                        NeeSample ns;
                        nee_prepare(s, key, b, k, rec.p, rec.n, out.mat_kind, out.W, ns);
                        n_sh++;
                        if (nee_is_void(ns))
                            continue;
                        n_sh_traced++;
                        TraceCounters t2{0, 0};
                        V3 c = nee_sample<true>(s, rp, key, b, k, rec.p, rec.n, out.mat_kind, out.W, &t2);
                        sh.boxes += t2.boxes, sh.prims += t2.prims;
                        sum += c / (float)rp.light_samples;
                    }
                if (!out.cont)
                    break;
                in.o = out.o, in.d = out.d, in.beta = out.beta, in.attenuation = out.attenuation;
                in.last_bsdf_pdf = out.last_bsdf_pdf;
            }
            if (!is_nan(sum.x) && !is_nan(sum.y) && !is_nan(sum.z))
                checksum += (double)sum.x + 2.0 * (double)sum.y + 3.0 * (double)sum.z;
        }
    printf("%s %dx%d x %d spp: nodes %d instances %d\n", argv[1], W, H, S, s.n_nodes, s.n_inst);
    printf("  extension rays %llu: %.3f boxes, %.3f prims per ray\n", n_ext, (double)ext.boxes / n_ext, (double)ext.prims / n_ext);
    printf("  shadow rays %llu (traced %llu): %.3f boxes, %.3f prims per traced ray\n", n_sh, n_sh_traced,
           (double)sh.boxes / (n_sh_traced ? n_sh_traced : 1), (double)sh.prims / (n_sh_traced ? n_sh_traced : 1));
    printf("  vertices per bounce: miss | material kinds 0..5 (b200pt.h B200PT_MAT_*)\n");
    for (int b = 0; b < 10; b++) {
        printf("   b%d:", b);
        for (int k = 0; k < 7; k++)
            printf(" %8llu", kinds[b][k]);
        printf("\n");
    }
    printf("  film checksum %.9e\n", checksum);
    return 0;
}
