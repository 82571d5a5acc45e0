// scene_compile.cpp -- scene JSON -> flat scene.  Re-statement of what build_scene does to the JSON
// (scene_parser.h:241-595) without building any object graph: every quirk that changes what gets
// rendered is kept and cited; the output is plain arrays for upload.
#include "flat_scene.h"
#include "bvh_build.h"

#include <cmath>
#include <cstdlib>
#include <cstring>
#include <exception>
#include <functional>
#include <map>
#include <random>
#include <thread>
#include <stdexcept>

namespace b200pt {

namespace {

const float kMauve[3] = {0.8f, 0.2f, 0.8f}; // scene_parser.h:16

struct Prim {
    bool valid = false;
    int kind = B200PT_PRIM_SPHERE;
    int plane = B200PT_PLANE_XZ;
    bool flip = false;
    float params[6] = {0, 0, 0, 0, 0, 0};
    int material = -1;           // surface material (for a medium: the boundary's, only used for light-ness)
    std::string material_type;   // wrapped_material::_type (scene_parser.h:57-68)
    bool medium = false;
    float density = 0.f;
    int phase_material = -1;
};

struct Builder {
    FlatScene out;
    std::string asset_dir;
    std::map<std::string, int> textures;
    std::map<std::string, std::pair<int, std::string>> materials; // id -> (index, type string)
    std::map<std::string, Prim> primitives;
    int last_id = 0; // generate_new_id (scene_parser.h:26-32)
    int error_texture = -1, error_material = -1;

    std::string new_id() { return std::to_string(last_id++); }

    int add_constant_texture(const float c[3], float alpha = 1.0f)
    {
        b200pt_texture t{};
        t.kind = B200PT_TEX_CONSTANT;
        t.color[0] = c[0];
        t.color[1] = c[1];
        t.color[2] = c[2];
        t.alpha = alpha;
        t.odd = t.even = -1;
        out.textures.push_back(t);
        return (int)out.textures.size() - 1;
    }

    int get_error_texture()
    {
        if (error_texture < 0)
            error_texture = add_constant_texture(kMauve);
        return error_texture;
    }

    int add_material(int kind, int texture, const float color[3] = nullptr, float ior = 1.45f, float power = 1.f,
                     bool two_sided = true)
    {
        b200pt_material m{};
        m.kind = kind;
        m.texture = texture;
        if (color) {
            m.color[0] = color[0];
            m.color[1] = color[1];
            m.color[2] = color[2];
        }
        m.ior = ior;
        m.power = power;
        m.two_sided = two_sided ? 1 : 0;
        out.materials.push_back(m);
        return (int)out.materials.size() - 1;
    }

    // error_material(): lambertian(MAUVE) (scene_parser.h:20-24)
    std::pair<int, std::string> get_error_material()
    {
        if (error_material < 0)
            error_material = add_material(B200PT_MAT_LAMBERTIAN, add_constant_texture(kMauve));
        return {error_material, "lambertian"};
    }
};

void json_vec3(const Json &j, float v[3])
{
    v[0] = j.at(0).as_float();
    v[1] = j.at(1).as_float();
    v[2] = j.at(2).as_float();
}

int lookup_texture(Builder &b, const std::string &id)
{
    auto it = b.textures.find(id);
    if (it == b.textures.end())
        throw std::runtime_error("scene: texture id '" + id + "' is not defined (the reference would dereference null)");
    return it->second;
}

// scene_parser.h:263-330
void parse_textures(Builder &b, const Json &scene)
{
    for (const Json &e : scene["textures"].items()) {
        if (e.value("skip", false))
            continue;
        if (!e.contains("id"))
            throw std::runtime_error("scene: texture without id (scene_parser.h:271 asserts)");
        std::string id = e["id"].as_string();
        int idx;
        if (!e.contains("data")) {
            idx = b.get_error_texture();
        } else {
            const Json &data = e["data"];
            std::string type = e["type"].as_string();
            if (type == "checker") {
                // checker_texture(even, odd, scale) texture.h:37
                int odd, even;
                float c[3];
                if (data["odd"].contains("texture")) {
                    odd = lookup_texture(b, data["odd"]["texture"].as_string());
                } else {
                    json_vec3(data["odd"]["color"], c);
                    odd = b.add_constant_texture(c);
                }
                if (data["even"].contains("texture")) {
                    even = lookup_texture(b, data["even"]["texture"].as_string());
                } else {
                    json_vec3(data["even"]["color"], c);
                    even = b.add_constant_texture(c);
                }
                b200pt_texture t{};
                t.kind = B200PT_TEX_CHECKER;
                t.alpha = 1.f;
                t.scale = data["scale"].as_float();
                t.odd = odd;
                t.even = even;
                b.out.textures.push_back(t);
                idx = (int)b.out.textures.size() - 1;
            } else if (type == "perlin") {
                b200pt_texture t{};
                t.kind = B200PT_TEX_PERLIN;
                t.alpha = 1.f;
                t.scale = (float)data.value("scale", 1.0);
                t.odd = t.even = -1;
                b.out.textures.push_back(t);
                idx = (int)b.out.textures.size() - 1;
            } else if (type == "png") {
                // decode_into_texture + from_4byte_vector (scene_parser.h:39-55, image.h:52-70)
                const std::string rel = data["path"].as_string();
                std::vector<std::string> tries = {rel};
                if (!b.asset_dir.empty() && !rel.empty() && rel[0] != '/') {
                    tries.push_back(b.asset_dir + "/" + rel);
                    tries.push_back(b.asset_dir + "/../" + rel);
                }
                std::vector<unsigned char> bytes;
                unsigned w = 0, h = 0;
                std::string err;
                bool ok = false;
                for (const std::string &path : tries) {
                    try {
                        decode_png_rgba8(path, bytes, w, h);
                        ok = true;
                        break;
                    } catch (const std::exception &ex) {
                        if (err.empty() || std::string(ex.what()).find("cannot open") == std::string::npos)
                            err = ex.what();
                    }
                }
                if (!ok) // the reference prints the decoder error and goes on with an empty image, which is UB
                    throw std::runtime_error("scene: " + err);
                ImageData img;
                img.width = (int32_t)w;
                img.height = (int32_t)h;
                img.rgba.resize(bytes.size());
                for (size_t k = 0; k < bytes.size(); k++)
                    img.rgba[k] = (float)(bytes[k] / 255.0); // `image[..] / 255.0` assigned to a float
                b.out.images.push_back(std::move(img));
                b200pt_texture t{};
                t.kind = B200PT_TEX_IMAGE;
                t.alpha = 1.f;
                t.odd = (int32_t)b.out.images.size() - 1;
                t.even = -1;
                b.out.textures.push_back(t);
                idx = (int)b.out.textures.size() - 1;
            } else {
                // "constant", and any unknown string: std::map::operator[] maps it to enum 0 = CONSTANT (scene.h:79)
                float c[3];
                json_vec3(data["color"], c);
                idx = b.add_constant_texture(c, (float)data.value("alpha", 1.0));
            }
        }
        b.textures.emplace(id, idx); // emplace: the first definition of an id wins
    }
}

// scene_parser.h:332-447
void parse_materials(Builder &b, const Json &scene)
{
    for (const Json &e : scene["materials"].items()) {
        if (e.value("skip", false))
            continue;
        if (!e.contains("id"))
            throw std::runtime_error("scene: material without id (scene_parser.h:338 asserts)");
        std::string id = e["id"].as_string();
        if (!e.contains("data")) {
            b.materials.emplace(id, b.get_error_material());
            continue;
        }
        const Json &data = e["data"];
        std::string type = e["type"].as_string();
        float c[3];
        if (type == "metal") {
            if (data.contains("color"))
                json_vec3(data["color"], c);
            else
                c[0] = c[1] = c[2] = 1.f;
            // fuzz (material.h:79-89) is stored and never read by the reference; it travels in `power` for the opt-in
            // corrected mode (device/pt_corrected.h)
            float fuzz = (float)data.value("roughness", 0.0);
            fuzz = fuzz < 1.f ? fuzz : 1.f;
            b.materials.emplace(id, std::make_pair(b.add_material(B200PT_MAT_METAL, -1, c, 1.45f, fuzz), std::string("metal")));
        } else if (type == "dielectric") {
            float ri = 1.450f;
            if (data.contains("ior"))
                ri = data["ior"].as_float();
            b.materials.emplace(id, std::make_pair(b.add_material(B200PT_MAT_DIELECTRIC, -1, nullptr, ri),
                                                   std::string("dielectric")));
        } else if (type == "diffuse_light") {
            float power = data.contains("power") ? data["power"].as_float() : 1.0f;
            bool two_sided = data.value("two_sided", true);
            int tex;
            if (data.contains("texture")) {
                tex = lookup_texture(b, data["texture"].as_string());
            } else {
                if (data.contains("color"))
                    json_vec3(data["color"], c);
                else
                    c[0] = c[1] = c[2] = 1.f;
                tex = b.add_constant_texture(c);
            }
            b.materials.emplace(id, std::make_pair(b.add_material(B200PT_MAT_DIFFUSE_LIGHT, tex, nullptr, 1.45f, power,
                                                                  two_sided),
                                                   std::string("diffuse_light")));
        } else if (type == "isotropic") {
            // enum ISOTROPIC has no case (scene_parser.h:348-446 `default: break`): the id stays undefined.
            b.out.warnings.push_back("material '" + id + "' of type isotropic is dropped, as in the reference");
        } else {
            // "lambertian", and any unknown string (maps to enum 0, scene.h:27)
            if (data.contains("color")) {
                json_vec3(data["color"], c);
                b.materials.emplace(id, std::make_pair(b.add_material(B200PT_MAT_LAMBERTIAN, b.add_constant_texture(c)),
                                                       std::string("lambertian")));
            } else if (data.contains("texture")) {
                int tex = lookup_texture(b, data["texture"].as_string());
                b.materials.emplace(id,
                                    std::make_pair(b.add_material(B200PT_MAT_LAMBERTIAN, tex), std::string("lambertian")));
            } else {
                b.materials.emplace(id, b.get_error_material());
            }
        }
    }
}

// parse_prim_or_instance (scene_parser.h:104-239)
Prim parse_primitive(Builder &b, const Json &e)
{
    Prim p;
    std::pair<int, std::string> mat;
    if (e.contains("material") && e["material"].contains("id")) {
        std::string mid = e["material"]["id"].as_string();
        auto it = b.materials.find(mid);
        if (it == b.materials.end())
            throw std::runtime_error("scene: material id '" + mid + "' is not defined (scene_parser.h:115 asserts)");
        mat = it->second;
    } else {
        b.out.warnings.push_back("WARNING! found misconfigured material for primitive");
        mat = b.get_error_material();
    }
    p.material = mat.first;
    p.material_type = mat.second;
    std::string type = e["type"].as_string();
    if (type == "sphere") {
        p.valid = true;
        p.kind = B200PT_PRIM_SPHERE;
        float r = 1.0f, o[3] = {0, 0, 0};
        if (e.contains("radius"))
            r = e["radius"].as_float();
        if (e.contains("origin"))
            json_vec3(e["origin"], o);
        p.params[0] = o[0];
        p.params[1] = o[1];
        p.params[2] = o[2];
        p.params[3] = r;
    } else if (type == "rect") {
        p.valid = true;
        p.kind = B200PT_PRIM_RECT;
        p.flip = e.value("flip", false);
        p.plane = B200PT_PLANE_XZ;
        if (e.contains("align")) {
            std::string a = e["align"].as_string();
            // plane_enum_mapping (primitive.h:18-25): unknown strings become enum 0 = XY
            p.plane = a == "xz" ? B200PT_PLANE_XZ : (a == "yz" ? B200PT_PLANE_YZ : B200PT_PLANE_XY);
        }
        if (e.contains("a0") && e.contains("b0") && e.contains("a1") && e.contains("b1")) {
            p.params[0] = e["a0"].as_float();
            p.params[1] = e["b0"].as_float();
            p.params[2] = e["a1"].as_float();
            p.params[3] = e["b1"].as_float();
            p.params[4] = e["c"].as_float();
        } else {
            float a = 1.f, bb = 1.f;
            if (e.contains("size")) {
                a = e["size"].at(0).as_float();
                bb = e["size"].at(1).as_float();
            }
            // rect(x, z, ...) primitive.h:132-136: (-x/2.0, -z/2.0, x/2.0, z/2.0, 0.0) -- double division, then float
            p.params[0] = (float)(-a / 2.0);
            p.params[1] = (float)(-bb / 2.0);
            p.params[2] = (float)(a / 2.0);
            p.params[3] = (float)(bb / 2.0);
            p.params[4] = 0.f;
        }
        if (!(p.params[0] < p.params[2]) || !(p.params[1] < p.params[3]))
            throw std::runtime_error("scene: rect needs x0<x1 and z0<z1 (primitive.h:147-148 asserts)");
    } else if (type == "box") {
        p.valid = true;
        p.kind = B200PT_PRIM_BOX;
        if (e.contains("p0") && e.contains("p1")) {
            json_vec3(e["p0"], p.params);
            json_vec3(e["p1"], p.params + 3);
        } else {
            float s[3] = {1, 1, 1};
            if (e.contains("size"))
                json_vec3(e["size"], s);
            // box(width,height,depth) primitive.h:230: +-size/2 in float
            for (int k = 0; k < 3; k++) {
                p.params[k] = -s[k] / 2;
                p.params[3 + k] = s[k] / 2;
            }
        }
        for (int k = 0; k < 3; k++)
            if (!(p.params[k] < p.params[3 + k]))
                throw std::runtime_error("scene: box needs p0<p1 on every axis (its rects assert, primitive.h:147-148)");
    } else if (type == "volume") {
        std::string pid = e["primitive"].as_string();
        float density = e["density"].as_float();
        float color[3] = {kMauve[0], kMauve[1], kMauve[2]};
        if (e.contains("color"))
            json_vec3(e["color"], color);
        auto it = b.primitives.find(pid);
        if (it == b.primitives.end() || !it->second.valid)
            throw std::runtime_error("scene: volume refers to undefined primitive '" + pid + "'");
        if (it->second.medium)
            throw std::runtime_error("scene: a volume whose boundary is itself a volume is not supported");
        p = it->second; // boundary geometry; the wrapped material stays the boundary's (scene_parser.h:231)
        p.medium = true;
        p.density = density;
        // constant_medium(b, d, vec3 a): phase_function = new isotropic(a) (volume.h:14-17)
        p.phase_material = b.add_material(B200PT_MAT_ISOTROPIC, b.add_constant_texture(color));
    } else {
        // "mesh" and unknown strings (enum 0 = MESH): nothing is built (scene_parser.h:126-128)
        p.valid = false;
    }
    return p;
}

} // namespace

void make_transform(const float scale[3], const float rotate[3], const float translate[3], float out[12])
{
    // AngleAxisf(rotate*M_PI, axis): the angle is a double product narrowed to float (transform3.h:20).
    float q[4] = {1.f, 0.f, 0.f, 0.f}; // w, x, y, z
    for (int axis = 0; axis < 3; axis++) {
        float angle = (float)((double)rotate[axis] * M_PI);
        float ha = 0.5f * angle;
        float b[4] = {cosf(ha), 0.f, 0.f, 0.f};
        b[1 + axis] = sinf(ha);
        float a[4] = {q[0], q[1], q[2], q[3]};
        q[0] = a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3];
        q[1] = a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2];
        q[2] = a[0] * b[2] + a[2] * b[0] + a[3] * b[1] - a[1] * b[3];
        q[3] = a[0] * b[3] + a[3] * b[0] + a[1] * b[2] - a[2] * b[1];
    }
    const float w = q[0], x = q[1], y = q[2], z = q[3];
    const float tx = 2.f * x, ty = 2.f * y, tz = 2.f * z;
    const float twx = tx * w, twy = ty * w, twz = tz * w;
    const float txx = tx * x, txy = ty * x, txz = tz * x;
    const float tyy = ty * y, tyz = tz * y, tzz = tz * z;
    float R[3][3];
    R[0][0] = 1.f - (tyy + tzz);
    R[0][1] = txy - twz;
    R[0][2] = txz + twy;
    R[1][0] = txy + twz;
    R[1][1] = 1.f - (txx + tzz);
    R[1][2] = tyz - twx;
    R[2][0] = txz - twy;
    R[2][1] = tyz + twx;
    R[2][2] = 1.f - (txx + tyy);
    for (int r = 0; r < 3; r++) {
        for (int c = 0; c < 3; c++)
            out[r * 4 + c] = R[r][c] * scale[c];
        out[r * 4 + 3] = translate[r];
    }
}

// Eigen 3.2.10's Affine3f::inverse() in the reference's own float arithmetic (Transform.h:1158-1185 ->
// LU/Inverse.h:120-159): 3x3 inverse by cofactors (det = c0*m00 + (c1*m10 + c2*m20), every entry = cofactor *
// (1/det)), translation = -(inv * t) summed left to right.  The reference re-derives this on every hit
// (transform3.h:51-54); bit-equal matrices are what keeps near-surface rays (|o - surface| << |o|) in agreement.
// Compiled without FMA contraction (x86-64 baseline), like the reference.
bool invert_affine(const float m[12], float out[12])
{
    auto M = [&](int r, int c) -> float { return m[r * 4 + c]; };
    auto cof = [&](int i, int j) -> float {
        const int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
        volatile float a = M(i1, j1) * M(i2, j2);
        volatile float b = M(i1, j2) * M(i2, j1);
        return a - b;
    };
    const float c0 = cof(0, 0), c1 = cof(1, 0), c2 = cof(2, 0);
    volatile float p0 = c0 * M(0, 0), p1 = c1 * M(1, 0), p2 = c2 * M(2, 0);
    volatile float s12 = p1 + p2;
    const float det = p0 + s12;
    if (det == 0.0f || !std::isfinite(det))
        return false;
    const float invdet = 1.0f / det;
    float inv[3][3];
    inv[0][0] = c0 * invdet;
    inv[0][1] = c1 * invdet;
    inv[0][2] = c2 * invdet;
    inv[1][0] = cof(0, 1) * invdet;
    inv[1][1] = cof(1, 1) * invdet;
    inv[1][2] = cof(2, 1) * invdet;
    inv[2][0] = cof(0, 2) * invdet;
    inv[2][1] = cof(1, 2) * invdet;
    inv[2][2] = cof(2, 2) * invdet;
    for (int r = 0; r < 3; r++) {
        for (int c = 0; c < 3; c++)
            out[r * 4 + c] = inv[r][c];
        volatile float q0 = inv[r][0] * m[3], q1 = inv[r][1] * m[7], q2 = inv[r][2] * m[11];
        volatile float q01 = q0 + q1;
        out[r * 4 + 3] = -(q01 + q2);
    }
    return true;
}

static void shuffle3(const float v[3], int plane, float o[3])
{
    // shuffle() primitive.h:104-121
    if (plane == B200PT_PLANE_XY) {
        o[0] = v[0];
        o[1] = v[2];
        o[2] = v[1];
    } else if (plane == B200PT_PLANE_YZ) {
        o[0] = v[1];
        o[1] = v[0];
        o[2] = v[2];
    } else {
        o[0] = v[0];
        o[1] = v[1];
        o[2] = v[2];
    }
}

Aabb instance_bounds(const b200pt_instance &inst)
{
    float lo[3], hi[3];
    const float *p = inst.params;
    if (inst.prim_kind == B200PT_PRIM_SPHERE) {
        for (int k = 0; k < 3; k++) {
            lo[k] = p[k] - p[3];
            hi[k] = p[k] + p[3];
        }
    } else if (inst.prim_kind == B200PT_PRIM_RECT) {
        // rect::bounding_box primitive.h:145-154: y -+ 0.001 is double arithmetic narrowed to float
        float a[3] = {p[0], (float)(p[4] - 0.001), p[1]};
        float b[3] = {p[2], (float)(p[4] + 0.001), p[3]};
        shuffle3(a, inst.plane, lo);
        shuffle3(b, inst.plane, hi);
    } else {
        // box: union of its six rect boxes (primitive.h:233-252, hittable_list.h:40-68)
        for (int k = 0; k < 3; k++) {
            lo[k] = (float)(p[k] - 0.001);
            hi[k] = (float)(p[3 + k] + 0.001);
        }
    }
    Aabb out;
    for (int k = 0; k < 3; k++) {
        out.lo[k] = 3.402823466e+38f;
        out.hi[k] = -3.402823466e+38f;
    }
    const float *m = inst.fwd;
    for (int i = 0; i < 2; i++)
        for (int j = 0; j < 2; j++)
            for (int k = 0; k < 2; k++) {
                float x = i ? hi[0] : lo[0], y = j ? hi[1] : lo[1], z = k ? hi[2] : lo[2];
                for (int c = 0; c < 3; c++) {
                    float v = m[c * 4 + 0] * x + m[c * 4 + 1] * y + m[c * 4 + 2] * z + m[c * 4 + 3];
                    if (v > out.hi[c])
                        out.hi[c] = v;
                    if (v < out.lo[c])
                        out.lo[c] = v;
                }
            }
    return out;
}

b200pt_camera make_camera(const float lookfrom[3], const float lookat[3], float vfov, float aspect, float aperture,
                          float focus_dist)
{
    b200pt_camera cam{};
    cam.lens_radius = aperture / 2;
    float theta = (float)(vfov * M_PI / 180);
    float half_height = tanf(theta / 2);
    float half_width = aspect * half_height;
    float w[3], u[3], v[3];
    float d[3] = {lookfrom[0] - lookat[0], lookfrom[1] - lookat[1], lookfrom[2] - lookat[2]};
    float len = sqrtf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
    for (int k = 0; k < 3; k++)
        w[k] = d[k] / len;
    const float vup[3] = {0.f, 1.f, 0.f}; // main.cpp:86
    float c[3] = {vup[1] * w[2] - vup[2] * w[1], vup[2] * w[0] - vup[0] * w[2], vup[0] * w[1] - vup[1] * w[0]};
    len = sqrtf(c[0] * c[0] + c[1] * c[1] + c[2] * c[2]);
    for (int k = 0; k < 3; k++)
        u[k] = c[k] / len;
    v[0] = w[1] * u[2] - w[2] * u[1];
    v[1] = w[2] * u[0] - w[0] * u[2];
    v[2] = w[0] * u[1] - w[1] * u[0];
    for (int k = 0; k < 3; k++) {
        cam.origin[k] = lookfrom[k];
        cam.u[k] = u[k];
        cam.v[k] = v[k];
        cam.w[k] = w[k];
        // camera.h:30-33
        cam.lower_left[k] =
            lookfrom[k] - (half_width * focus_dist) * u[k] - (half_height * focus_dist) * v[k] - focus_dist * w[k];
        cam.horizontal[k] = (2 * half_width * focus_dist) * u[k];
        cam.vertical[k] = (2 * half_height * focus_dist) * v[k];
    }
    return cam;
}

namespace {

struct RefRng {
    // random.h:9-15: uniform_real_distribution<double>(0,1) bound to a default-constructed mt19937
    std::uniform_real_distribution<double> distribution{0.0, 1.0};
    std::mt19937 generator;
    double operator()() { return distribution(generator); }
};

struct SortCtx {
    const Aabb *boxes;
    int axis;
};

// box_{x,y,z}_compare (bvh.h:71-131): never returns 0
int ref_box_compare(const void *a, const void *b, void *ctx)
{
    const SortCtx &c = *(const SortCtx *)ctx;
    long ia = *(const long *)a, ib = *(const long *)b;
    float d = c.boxes[ia].lo[c.axis] - c.boxes[ib].lo[c.axis];
    return d < 0.0 ? -1 : 1;
}

void ref_bvh_walk(RefRng &rng, const std::vector<Aabb> &boxes, long *l, int n, std::vector<int> &dup)
{
    int axis = int(3 * rng()); // bvh.h:135
    SortCtx ctx{boxes.data(), axis > 2 ? 2 : axis};
    // glibc's qsort IS qsort_r with a null argument: same algorithm, same element size (a pointer) as bvh.h:137-148,
    // hence the same permutation -- ties included, which the never-zero comparator leaves to the algorithm
    qsort_r(l, n, sizeof(long), ref_box_compare, &ctx);
    if (n == 1) {
        dup[l[0]] = 1; // left = right = l[0] (bvh.h:150-153)
    } else if (n > 2) {
        ref_bvh_walk(rng, boxes, l, n / 2, dup);
        ref_bvh_walk(rng, boxes, l + n / 2, n - n / 2, dup);
    }
}

} // namespace

void replay_reference_rng(FlatScene &scene)
{
    RefRng rng;
    // perlin_generate (texture.h:91-102)
    scene.perlin_ranvec.assign(256 * 3, 0.f);
    for (int i = 0; i < 256; i++) {
        double x = 2 * rng() - 1, y = 2 * rng() - 1, z = 2 * rng() - 1;
        float fx = (float)x, fy = (float)y, fz = (float)z;
        float len = sqrtf(fx * fx + fy * fy + fz * fz);
        scene.perlin_ranvec[i * 3 + 0] = fx / len;
        scene.perlin_ranvec[i * 3 + 1] = fy / len;
        scene.perlin_ranvec[i * 3 + 2] = fz / len;
    }
    // perlin_generate_perm x3 (texture.h:67-89, 181-183)
    scene.perlin_perm.assign(3 * 256, 0);
    for (int t = 0; t < 3; t++) {
        int32_t *p = &scene.perlin_perm[t * 256];
        for (int i = 0; i < 256; i++)
            p[i] = i;
        for (int i = 255; i > 0; i--) {
            int target = int(rng() * (i + 1));
            int32_t tmp = p[i];
            p[i] = p[target];
            p[target] = tmp;
        }
    }
    // bvh_node::bvh_node over the instance list in build_scene order (scene_parser.h:594)
    const int n = (int)scene.instances.size();
    std::vector<Aabb> boxes(n);
    std::vector<long> order(n);
    std::vector<int> dup(n, 0);
    for (int i = 0; i < n; i++) {
        boxes[i] = instance_bounds(scene.instances[i]);
        order[i] = i;
    }
    // the tree the kernels walk does not depend on the order of the instance table, only its leaf labels do: for a
    // large scene it is built on a second thread meanwhile, over the build-order boxes, and relabelled below
    std::shared_ptr<BuiltBvh> built;
    std::exception_ptr build_error;
    std::thread builder;
    if (n >= 1024)
        builder = std::thread([&] {
            try {
                built = std::make_shared<BuiltBvh>(build_bvh(boxes));
            } catch (...) {
                build_error = std::current_exception();
            }
        });
    if (n > 0)
        ref_bvh_walk(rng, boxes, order.data(), n, dup); // qsort and recursion only: does not throw
    if (builder.joinable())
        builder.join();
    if (build_error)
        std::rethrow_exception(build_error);
    // `order` now lists the instances as the leaves of the reference tree, left to right.  bvh_node::hit keeps the
    // RIGHT child's record when both children report the same t (bvh.h:57-60), so among exactly coincident
    // surfaces (a box standing on the floor) the right-most leaf wins.  The instance table is stored in that
    // order, which turns the rule into "the higher instance index wins a tie" (trace_closest).
    std::vector<b200pt_instance> sorted(n);
    std::vector<int32_t> new_index(n);
    for (int i = 0; i < n; i++) {
        sorted[i] = scene.instances[order[i]];
        if (dup[order[i]])
            sorted[i].flags |= B200PT_INST_BVH_DUP;
        else
            sorted[i].flags &= ~B200PT_INST_BVH_DUP;
        new_index[order[i]] = i;
    }
    scene.instances.swap(sorted);
    for (int32_t &l : scene.lights)
        l = new_index[l];
    if (built) {
        auto relabel = [&](int32_t &ref) {
            if (ref < 0)
                ref = ~new_index[~ref];
        };
        for (pt::PtNode &nd : built->nodes) {
            relabel(nd.ref0);
            relabel(nd.ref1);
        }
        for (pt::PtNode4 &nd : built->nodes4)
            for (int k = 0; k < 4; k++)
                relabel(nd.ref[k]);
        scene.prebuilt_bvh = built;
    }
}

FlatScene compile_scene(const Json &scene, int film_width, int film_height, const std::string &asset_dir)
{
    Builder b;
    b.asset_dir = asset_dir;
    parse_textures(b, scene);
    parse_materials(b, scene);

    // primitives: NB no "skip" test here (scene_parser.h:450-462)
    for (const Json &e : scene["primitives"].items()) {
        std::string id = e.contains("id") ? e["id"].as_string() : b.new_id();
        Prim p = parse_primitive(b, e);
        b.primitives.emplace(id, p);
    }
    // "direct" instances are turned into anonymous primitives first, skipped or not (scene_parser.h:464-480)
    std::vector<std::string> instance_prim_ids;
    for (const Json &e : scene["instances"].items()) {
        if (e["type"].as_string() == "ref") {
            instance_prim_ids.push_back(e["primitive"].contains("id") ? e["primitive"]["id"].as_string() : "");
            continue;
        }
        std::string id = b.new_id();
        b.primitives.emplace(id, parse_primitive(b, e["primitive"]));
        instance_prim_ids.push_back(id);
    }
    size_t k = 0;
    for (const Json &e : scene["instances"].items()) {
        const std::string &pid = instance_prim_ids[k++];
        if (e.value("skip", false))
            continue;
        float scale[3] = {1, 1, 1}, rotate[3] = {0, 0, 0}, translate[3] = {0, 0, 0};
        if (e.contains("transform")) {
            const Json &t = e["transform"];
            if (t.contains("scale") && t["scale"].is_array()) {
                json_vec3(t["scale"], scale);
            } else {
                float f = (float)t.value("scale", 1.0);
                scale[0] = scale[1] = scale[2] = f;
            }
            if (t.contains("rotate"))
                json_vec3(t["rotate"], rotate);
            if (t.contains("translate"))
                json_vec3(t["translate"], translate);
        }
        auto it = b.primitives.find(pid);
        if (it == b.primitives.end())
            throw std::runtime_error("scene: instance refers to undefined primitive '" + pid +
                                     "' (scene_parser.h:534 asserts)");
        const Prim &p = it->second;
        if (!p.valid)
            throw std::runtime_error("scene: instance of a primitive that builds nothing (mesh/unknown type)");
        b200pt_instance inst{};
        make_transform(scale, rotate, translate, inst.fwd);
        inst.prim_kind = p.kind;
        inst.plane = p.plane;
        inst.flags = (p.medium ? B200PT_INST_MEDIUM : 0u) | (p.flip ? B200PT_INST_RECT_FLIP : 0u);
        inst.material = p.medium ? p.phase_material : p.material;
        memcpy(inst.params, p.params, sizeof(inst.params));
        inst.density = p.density;
        inst.light_index = -1;
        if (p.material_type == "diffuse_light") { // scene_parser.h:541-549
            inst.light_index = (int)b.out.lights.size();
            b.out.lights.push_back((int)b.out.instances.size());
        }
        b.out.instances.push_back(inst);
    }
    if (b.out.instances.empty())
        throw std::runtime_error("scene: no instances");

    // background (scene_parser.h:552-577)
    if (scene.contains("world") && scene["world"].contains("texture")) {
        b.out.background = lookup_texture(b, scene["world"]["texture"].as_string());
    } else if (scene.contains("world") && scene["world"].contains("color")) {
        float c[3];
        json_vec3(scene["world"]["color"], c);
        b.out.background = b.add_constant_texture(c);
    } else {
        b.out.background = b.add_constant_texture(kMauve);
    }

    // setup_camera (main.cpp:86-104)
    const Json &cj = scene["camera"];
    float from[3], at[3];
    json_vec3(cj["look_from"], from);
    json_vec3(cj["look_at"], at);
    float vfov = (float)cj.value("fov", 30.0);
    float aperture = (float)cj.value("aperture", 0.0);
    float focus = (float)cj.value("dist_to_focus", 10.0);
    b.out.camera = make_camera(from, at, vfov, float(film_width) / float(film_height), aperture, focus);

    replay_reference_rng(b.out);
    return b.out;
}

b200pt_scene_desc FlatScene::desc() const
{
    b200pt_scene_desc d{};
    d.instances = instances.data();
    d.n_instances = (int32_t)instances.size();
    d.materials = materials.data();
    d.n_materials = (int32_t)materials.size();
    d.textures = textures.data();
    d.n_textures = (int32_t)textures.size();
    d.lights = lights.data();
    d.n_lights = (int32_t)lights.size();
    d.background = background;
    d.camera = camera;
    d.perlin_ranvec = perlin_ranvec.empty() ? nullptr : perlin_ranvec.data();
    d.perlin_perm = perlin_perm.empty() ? nullptr : perlin_perm.data();
    image_views.clear();
    for (const ImageData &im : images)
        image_views.push_back(b200pt_image{im.width, im.height, im.rgba.data()});
    d.images = image_views.empty() ? nullptr : image_views.data();
    d.n_images = (int32_t)image_views.size();
    return d;
}

FlatScene FlatScene::from_desc(const b200pt_scene_desc &d)
{
    FlatScene s;
    if (!d.instances || d.n_instances <= 0)
        throw std::runtime_error("scene desc: no instances");
    s.instances.assign(d.instances, d.instances + d.n_instances);
    if (d.materials && d.n_materials > 0)
        s.materials.assign(d.materials, d.materials + d.n_materials);
    if (d.textures && d.n_textures > 0)
        s.textures.assign(d.textures, d.textures + d.n_textures);
    if (d.lights && d.n_lights > 0)
        s.lights.assign(d.lights, d.lights + d.n_lights);
    s.background = d.background;
    s.camera = d.camera;
    if (d.perlin_ranvec)
        s.perlin_ranvec.assign(d.perlin_ranvec, d.perlin_ranvec + 256 * 3);
    if (d.perlin_perm)
        s.perlin_perm.assign(d.perlin_perm, d.perlin_perm + 3 * 256);
    for (int k = 0; d.images && k < d.n_images; k++) {
        const b200pt_image &im = d.images[k];
        if (im.width <= 0 || im.height <= 0 || !im.rgba)
            throw std::runtime_error("scene desc: bad image");
        ImageData img;
        img.width = im.width;
        img.height = im.height;
        img.rgba.assign(im.rgba, im.rgba + (size_t)4 * im.width * im.height);
        s.images.push_back(std::move(img));
    }
    for (const auto &t : s.textures)
        if (t.kind == B200PT_TEX_IMAGE && (t.odd < 0 || t.odd >= (int)s.images.size()))
            throw std::runtime_error("scene desc: image texture refers to an image out of range");
    for (const auto &i : s.instances) {
        if (i.material < 0 || i.material >= (int)s.materials.size())
            throw std::runtime_error("scene desc: instance material out of range");
        if (i.prim_kind < 0 || i.prim_kind > B200PT_PRIM_BOX)
            throw std::runtime_error("scene desc: bad prim_kind");
    }
    for (const auto &m : s.materials) {
        bool needs_tex = m.kind != B200PT_MAT_METAL && m.kind != B200PT_MAT_DIELECTRIC;
        if (needs_tex && (m.texture < 0 || m.texture >= (int)s.textures.size()))
            throw std::runtime_error("scene desc: material texture out of range");
    }
    for (int l : s.lights)
        if (l < 0 || l >= (int)s.instances.size())
            throw std::runtime_error("scene desc: light id out of range");
    if (s.background < 0 || s.background >= (int)s.textures.size())
        throw std::runtime_error("scene desc: background texture out of range");
    return s;
}

} // namespace b200pt
