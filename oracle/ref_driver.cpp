// oracle/ref_driver.cpp -- TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// Harness around the UNMODIFIED reference (gillett-hernandez/pathtrace).  The reference is a
// single translation unit (main.cpp + headers); we include it where it lies (-I<reference root>),
// renaming its main(), and drive its own factories:
//   build_scene      scene_parser.h:241     setup_camera          main.cpp:86
//   World::hit       world.h:17             integrator_from_config main.cpp:30
//   NEEIterative::color integrator.h:176    RecursivePT::color    integrator.h:29
// No reference source is copied into this repository; the binary is written to oracle/_ref/.
//
// Modes (all single-threaded and therefore bit-reproducible -- the reference's global mt19937
// has a fixed default seed, random.h:9-15; `burn` discards K draws first to get another replica):
//   scene  <scene.json> <film_w> <film_h> <out.json>      leaves of the BVH, lights, camera
//   hit    <scene.json> <rays.f32> <out.bin> [burn]       closest hit per ray (n x 8 float32 in)
//   color  <config.json> <out.f64> <nsamples> <clamp> [burn]   per-pixel sample statistics
//   main   (cwd holds config.json)                        the reference's own main(), untouched
//   film   <fb.f32> <w> <h> <n_samples> <exposure> <out.ppm> <out.json>
//                                                         calculate_luminance (helpers.h:146-168) + output_to_file
//                                                         (renderer.h:24-55) on a supplied framebuffer of sums
//   project <scene.json> <film_w> <film_h> <paths.txt> <out3d.txt> <out2d.txt> <out.json>
//                                                         camera::project (camera.h:49-109) per point and
//                                                         handle_traced_paths (renderer.h:57-96) on supplied paths
#define main ref_main
#include "main.cpp"
#undef main

#include <cstdio>
#include <cstring>
#include <typeinfo>
#include <set>

static void burn_draws(long k)
{
    for (long i = 0; i < k; i++)
        random_double();
}

static json read_json(const char *path)
{
    std::ifstream f(path);
    if (!f.good())
    {
        fprintf(stderr, "ref_driver: cannot open %s\n", path);
        exit(2);
    }
    json j;
    f >> j;
    return j;
}

// build_scene echoes every element to stdout; silence it.
struct quiet_stdout
{
    std::streambuf *saved;
    std::ostringstream sink;
    quiet_stdout() { saved = std::cout.rdbuf(sink.rdbuf()); }
    ~quiet_stdout() { std::cout.rdbuf(saved); }
};

static void collect_leaves(hittable *h, std::vector<instance *> &leaves, std::set<hittable *> &seen, std::vector<int> &dup)
{
    if (bvh_node *n = dynamic_cast<bvh_node *>(h))
    {
        if (n->left == n->right && dynamic_cast<instance *>(n->left))
        {
            // n==1 leaf (bvh.h:150-153): primitive is intersected twice
            instance *i = dynamic_cast<instance *>(n->left);
            if (!seen.count(i))
            {
                seen.insert(i);
                leaves.push_back(i);
                dup.push_back(1);
            }
            return;
        }
        collect_leaves(n->left, leaves, seen, dup);
        collect_leaves(n->right, leaves, seen, dup);
        return;
    }
    instance *i = dynamic_cast<instance *>(h);
    if (i && !seen.count(i))
    {
        seen.insert(i);
        leaves.push_back(i);
        dup.push_back(0);
    }
}

static json vec_json(const vec3 &v) { return json::array({v.x(), v.y(), v.z()}); }

static json texture_json(texture *t)
{
    json j;
    if (constant_texture *c = dynamic_cast<constant_texture *>(t))
    {
        j["type"] = "constant";
        j["color"] = vec_json(c->color);
        j["alpha"] = c->a;
    }
    else if (checker_texture *c = dynamic_cast<checker_texture *>(t))
    {
        j["type"] = "checker";
        j["scale"] = c->scale;
        j["odd"] = texture_json(c->odd);
        j["even"] = texture_json(c->even);
    }
    else if (noise_texture *c = dynamic_cast<noise_texture *>(t))
    {
        j["type"] = "perlin";
        j["scale"] = c->scale;
    }
    else
    {
        j["type"] = "other";
    }
    return j;
}

static json material_json(material *m)
{
    json j;
    if (lambertian *l = dynamic_cast<lambertian *>(m))
    {
        j["type"] = "lambertian";
        j["texture"] = texture_json(l->albedo);
    }
    else if (metal *l = dynamic_cast<metal *>(m))
    {
        j["type"] = "metal";
        j["color"] = vec_json(l->albedo);
        j["fuzz"] = l->fuzz;
    }
    else if (dielectric *l = dynamic_cast<dielectric *>(m))
    {
        j["type"] = "dielectric";
        j["ior"] = l->ref_idx;
    }
    else if (diffuse_light *l = dynamic_cast<diffuse_light *>(m))
    {
        j["type"] = "diffuse_light";
        j["texture"] = texture_json(l->emit);
        j["power"] = l->power;
        j["two_sided"] = l->two_sided;
    }
    else if (isotropic *l = dynamic_cast<isotropic *>(m))
    {
        j["type"] = "isotropic";
        j["texture"] = texture_json(l->albedo);
    }
    else
    {
        j["type"] = "null";
    }
    return j;
}

static json prim_json(hittable *h)
{
    json j;
    if (sphere *s = dynamic_cast<sphere *>(h))
    {
        j["type"] = "sphere";
        j["center"] = vec_json(s->center);
        j["radius"] = s->radius;
        j["material"] = material_json(s->mat_ptr);
    }
    else if (rect *r = dynamic_cast<rect *>(h))
    {
        j["type"] = "rect";
        j["x0"] = r->x0;
        j["z0"] = r->z0;
        j["x1"] = r->x1;
        j["z1"] = r->z1;
        j["y"] = r->y;
        j["plane"] = (int)r->type;
        j["normal"] = r->normal;
        j["material"] = material_json(r->mp);
    }
    else if (box *b = dynamic_cast<box *>(h))
    {
        j["type"] = "box";
        j["p0"] = vec_json(b->p0);
        j["p1"] = vec_json(b->p1);
        hittable_list *g = dynamic_cast<hittable_list *>(b->group);
        j["material"] = material_json(dynamic_cast<rect *>(g->list[0])->mp);
    }
    else if (constant_medium *m = dynamic_cast<constant_medium *>(h))
    {
        j["type"] = "volume";
        j["density"] = m->density;
        j["boundary"] = prim_json(m->boundary);
        j["material"] = material_json(m->phase_function);
    }
    else
    {
        j["type"] = "other";
    }
    return j;
}

struct loaded_scene
{
    World *world;
    std::vector<instance *> leaves;
    std::vector<int> dup;
    std::map<hittable *, int> index;
};

static loaded_scene load_scene(const char *scene_path)
{
    loaded_scene ls;
    json scene = read_json(scene_path);
    {
        quiet_stdout q;
        ls.world = build_scene(scene);
    }
    std::set<hittable *> seen;
    collect_leaves(ls.world->ptr, ls.leaves, seen, ls.dup);
    for (size_t i = 0; i < ls.leaves.size(); i++)
        ls.index[ls.leaves[i]] = (int)i;
    return ls;
}

static int mode_scene(int argc, char **argv)
{
    if (argc < 6)
        return 2;
    loaded_scene ls = load_scene(argv[2]);
    int fw = atoi(argv[3]), fh = atoi(argv[4]);
    json scene = read_json(argv[2]);
    camera cam = setup_camera(scene["camera"], float(fw) / float(fh));
    json out;
    out["n_leaves"] = ls.leaves.size();
    json jl = json::array();
    for (size_t i = 0; i < ls.leaves.size(); i++)
    {
        instance *in = ls.leaves[i];
        json e;
        Eigen::Matrix4f m = in->transform._transform.matrix();
        json mj = json::array();
        for (int r = 0; r < 3; r++)
            for (int c = 0; c < 4; c++)
                mj.push_back(m(r, c));
        e["transform"] = mj;
        e["bbox_min"] = vec_json(in->bbox.min());
        e["bbox_max"] = vec_json(in->bbox.max());
        e["prim"] = prim_json(in->ptr);
        e["bvh_duplicated"] = ls.dup[i];
        int light_idx = -1;
        for (size_t l = 0; l < ls.world->lights.size(); l++)
            if (ls.world->lights[l] == in)
                light_idx = (int)l;
        e["light_index"] = light_idx;
        jl.push_back(e);
    }
    out["leaves"] = jl;
    out["n_lights"] = ls.world->lights.size();
    out["background"] = texture_json(ls.world->background);
    json jc;
    jc["origin"] = vec_json(cam.origin);
    jc["lower_left_corner"] = vec_json(cam.lower_left_corner);
    jc["horizontal"] = vec_json(cam.horizontal);
    jc["vertical"] = vec_json(cam.vertical);
    jc["u"] = vec_json(cam.u);
    jc["v"] = vec_json(cam.v);
    jc["w"] = vec_json(cam.w);
    jc["lens_radius"] = cam.lens_radius;
    out["camera"] = jc;
    // Perlin tables (texture.h:180-183) are drawn from the global RNG at static-init time.
    json pj;
    json rv = json::array();
    for (int i = 0; i < 256; i++)
        rv.push_back(vec_json(perlin::ranvec[i]));
    pj["ranvec"] = rv;
    json px = json::array(), py = json::array(), pz = json::array();
    for (int i = 0; i < 256; i++)
    {
        px.push_back(perlin::perm_x[i]);
        py.push_back(perlin::perm_y[i]);
        pz.push_back(perlin::perm_z[i]);
    }
    pj["perm_x"] = px;
    pj["perm_y"] = py;
    pj["perm_z"] = pz;
    out["perlin"] = pj;
    std::ofstream of(argv[5]);
    of << std::setprecision(9) << out.dump(1) << std::endl;
    return 0;
}

struct hit_out
{
    int32_t id;
    float t, p[3], n[3], u, v;
};

static int mode_hit(int argc, char **argv)
{
    if (argc < 5)
        return 2;
    loaded_scene ls = load_scene(argv[2]);
    if (argc > 5)
        burn_draws(atol(argv[5]));
    FILE *f = fopen(argv[3], "rb");
    if (!f)
        return 2;
    fseek(f, 0, SEEK_END);
    long bytes = ftell(f);
    fseek(f, 0, SEEK_SET);
    size_t n = bytes / (8 * sizeof(float));
    std::vector<float> rays(n * 8);
    if (fread(rays.data(), sizeof(float), n * 8, f) != n * 8)
        return 2;
    fclose(f);
    std::vector<hit_out> out(n);
    for (size_t i = 0; i < n; i++)
    {
        const float *q = &rays[i * 8];
        ray r(vec3(q[0], q[1], q[2]), vec3(q[3], q[4], q[5]), 0.0f);
        hit_record rec;
        rec.u = 0;
        rec.v = 0;
        hit_out &o = out[i];
        memset(&o, 0, sizeof(o));
        if (ls.world->hit(r, q[6], q[7], rec))
        {
            o.id = ls.index.count(rec.primitive) ? ls.index[rec.primitive] : -2;
            o.t = rec.t;
            for (int c = 0; c < 3; c++)
            {
                o.p[c] = rec.p[c];
                o.n[c] = rec.normal[c];
            }
            o.u = rec.u;
            o.v = rec.v;
        }
        else
        {
            o.id = -1;
        }
    }
    FILE *g = fopen(argv[4], "wb");
    fwrite(out.data(), sizeof(hit_out), n, g);
    fclose(g);
    return 0;
}

// Per-pixel sample statistics of integrator->color, sampled exactly as Tiled::compute does
// (renderer.h:648-670): jittered (u,v), cam.get_ray, one draw for the trace-probability test.
// Output (float64): header {W,H,N,clamp,total_rays}, then per pixel (row j=0 is the image BOTTOM,
// as in the reference framebuffer): mean[3] (unclamped), cmean[3], cM2[3] (clamped to [-clamp, clamp]).
static int mode_color(int argc, char **argv)
{
    if (argc < 6)
        return 2;
    json jconfig = read_json(argv[2]);
    Config config(jconfig);
    int N = atoi(argv[4]);
    double clampv = atof(argv[5]);
    if (argc > 6)
        burn_draws(atol(argv[6]));
    json scene = read_json(config.scene_path.c_str());
    World *world;
    Integrator *integrator;
    {
        quiet_stdout q;
        world = build_scene(scene);
        world->config = config;
        integrator = integrator_from_config(world, config);
    }
    int W = config.film.width, H = config.film.height;
    camera cam = setup_camera(scene["camera"], float(W) / float(H));
    std::vector<double> out(5 + (size_t)W * H * 9, 0.0);
    long total = 0;
    for (int j = 0; j < H; j++)
    {
        for (int i = 0; i < W; i++)
        {
            double mean[3] = {0, 0, 0}, cmean[3] = {0, 0, 0}, cm2[3] = {0, 0, 0};
            for (int s = 0; s < N; s++)
            {
                long count = 0;
                float u = float(i + random_double()) / float(W);
                float v = float(j + random_double()) / float(H);
                ray r = cam.get_ray(u, v);
                random_double(); // trace_probability test, renderer.h:652
                vec3 col = de_nan(integrator->color(r, 0, &count, nullptr));
                total += count;
                for (int c = 0; c < 3; c++)
                {
                    double x = col[c];
                    mean[c] += (x - mean[c]) / (s + 1);
                    double xc = x > clampv ? clampv : (x < -clampv ? -clampv : x); // two-sided: a Perlin albedo can be negative (texture.h:191)
                    double d = xc - cmean[c];
                    cmean[c] += d / (s + 1);
                    cm2[c] += d * (xc - cmean[c]);
                }
            }
            double *o = &out[5 + ((size_t)j * W + i) * 9];
            for (int c = 0; c < 3; c++)
            {
                o[c] = mean[c];
                o[3 + c] = cmean[c];
                o[6 + c] = cm2[c];
            }
        }
    }
    out[0] = W;
    out[1] = H;
    out[2] = N;
    out[3] = clampv;
    out[4] = (double)total;
    FILE *g = fopen(argv[3], "wb");
    fwrite(out.data(), sizeof(double), out.size(), g);
    fclose(g);
    return 0;
}

// The reference's film path on a framebuffer handed in by the test: fb.f32 holds H x W x 3 float32 SUMS, row 0 =
// image bottom (the layout of Renderer::framebuffer, renderer.h:682).
static int mode_film(int argc, char **argv)
{
    if (argc < 9)
        return 2;
    const int W = atoi(argv[3]), H = atoi(argv[4]), n = atoi(argv[5]);
    const float exposure = (float)atof(argv[6]);
    std::vector<float> raw((size_t)W * H * 3);
    FILE *f = fopen(argv[2], "rb");
    if (!f || fread(raw.data(), sizeof(float), raw.size(), f) != raw.size())
        return 2;
    fclose(f);
    vec3 **fb = new vec3 *[H];
    for (int j = 0; j < H; j++)
    {
        fb[j] = new vec3[W];
        for (int i = 0; i < W; i++)
        {
            const float *q = &raw[((size_t)j * W + i) * 3];
            fb[j][i] = vec3(q[0], q[1], q[2]);
        }
    }
    float max_luminance = 0, total_luminance = 0, avg_luminance = 0;
    calculate_luminance(fb, W, H, n, (long)W * H, max_luminance, total_luminance, avg_luminance);
    auto out = std::make_shared<std::ofstream>(argv[7], std::ios::binary);
    output_to_file(out, fb, W, H, n, max_luminance, exposure, 0.0f);
    out->close();
    json j;
    j["max_luminance"] = max_luminance;
    j["total_luminance"] = total_luminance;
    j["avg_luminance"] = avg_luminance;
    std::ofstream oj(argv[8]);
    oj << std::setprecision(9) << j.dump() << std::endl;
    return 0;
}

// paths.txt: one "x y z" per line, a blank line ends a path (an empty path is allowed: two blank lines).
static int mode_project(int argc, char **argv)
{
    if (argc < 9)
        return 2;
    json scene = read_json(argv[2]);
    const int W = atoi(argv[3]), H = atoi(argv[4]);
    camera cam = setup_camera(scene["camera"], float(W) / float(H));
    paths all;
    path *cur = new path();
    std::ifstream in(argv[5]);
    std::string line;
    json pts = json::array();
    while (std::getline(in, line))
    {
        if (line.empty())
        {
            all.push_back(cur);
            cur = new path();
            continue;
        }
        float x, y, z;
        if (sscanf(line.c_str(), "%g %g %g", &x, &y, &z) != 3)
            return 2;
        cur->push_back(vec3(x, y, z));
        float px = 0, py = 0;
        bool ok = cam.project(vec3(x, y, z), px, py);
        pts.push_back(json::array({px, py, ok ? 1 : 0}));
    }
    if (!cur->empty())
        all.push_back(cur);
    int added = 0;
    handle_traced_paths(argv[6], argv[7], cam, added, 1, &all);
    json j;
    j["added_paths"] = added;
    j["projected"] = pts;
    std::ofstream oj(argv[8]);
    oj << std::setprecision(9) << j.dump() << std::endl;
    return 0;
}

int main(int argc, char **argv)
{
    if (argc < 2)
    {
        fprintf(stderr, "usage: ref_driver scene|hit|color|main|film|project ...\n");
        return 2;
    }
    if (!strcmp(argv[1], "scene"))
        return mode_scene(argc, argv);
    if (!strcmp(argv[1], "hit"))
        return mode_hit(argc, argv);
    if (!strcmp(argv[1], "color"))
        return mode_color(argc, argv);
    if (!strcmp(argv[1], "main"))
        return ref_main(argc, argv);
    if (!strcmp(argv[1], "film"))
        return mode_film(argc, argv);
    if (!strcmp(argv[1], "project"))
        return mode_project(argc, argv);
    return 2;
}
