#!/usr/bin/env python3
"""Writes the benchmark scenes in the reference's scene-JSON schema (scene_parser.h:241-595).

The four BASELINE.json configs name scenes by the reference's file names; the data below restates
their content (camera, materials, primitives, instances) with small builder helpers so that
`/root/reference` is not needed at run time.  tests/test_scenes.py checks, when the reference is
mounted, that every file written here is semantically identical (json.load equality) to the
reference's own scenes/<name>.json.  `book2_final` is the synthetic config 5.

usage: python scenes/make_scenes.py [outdir]
"""
import json
import os
import sys

import numpy as np


def lambertian(id_, rgb):
    return {"id": id_, "type": "lambertian", "data": {"color": list(rgb)}}


def light(id_, rgb):
    return {"id": id_, "type": "diffuse_light", "data": {"color": list(rgb)}}


def ref(prim_id, **transform):
    inst = {"type": "ref", "primitive": {"id": prim_id}}
    if transform:
        inst["transform"] = transform
    return inst


def direct(prim, skip=False, **transform):
    inst = {}
    if skip:
        inst["skip"] = True
    inst["type"] = "direct"
    inst["primitive"] = prim
    if transform:
        inst["transform"] = transform
    return inst


def rect(material, size, **extra):
    p = {"type": "rect", "material": {"id": material}, "size": list(size)}
    p.update(extra)
    return p


def sphere(material, radius=None):
    p = {"type": "sphere", "material": {"id": material}}
    if radius is not None:
        p["radius"] = radius
    return p


def box(material, size):
    return {"type": "box", "material": {"id": material}, "size": list(size)}


def camera(look_from, look_at, fov, focus):
    return {"look_from": look_from, "look_at": look_at, "fov": fov, "aperture": 0.0, "dist_to_focus": focus}


def scene(cam, world_rgb, materials, primitives, instances, textures=()):
    return {"camera": cam, "world": {"color": world_rgb}, "assets": [], "textures": list(textures),
            "materials": materials, "primitives": primitives, "instances": instances}


SKIPPED_SPHERE_LIGHT = direct(sphere("light"), skip=True, scale=[100.0, 20.0, 100.0], translate=[273, 200, 171])


def cornell_shell():
    """Floor, ceiling, back wall (refs to one 555x555 rect), green and red side walls."""
    return [
        ref("white_wall", translate=[277.5, 0.0, 277.5]),
        ref("white_wall", rotate=[1.0, 0.0, 0.0], translate=[277.5, 555, 277.5]),
        ref("white_wall", rotate=[1.5, 0, 0], translate=[277.5, 277.5, 555]),
        direct(rect("green", [555, 555], align="yz", flip=True), translate=[555, 277.5, 277.5]),
        direct(rect("red", [555, 555], align="yz"), translate=[0, 277.5, 277.5]),
    ]


TALL_BOX = direct(box("white", [165, 330, 165]), translate=[347.5, 165, 377.5], rotate=[0.0, 0.05, 0.0])
WHITE_WALL = dict(id="white_wall", **rect("white", [555, 555]))


def three_orbs():
    mats = [lambertian("green", [0.12, 0.45, 0.15]), lambertian("white", [0.7, 0.7, 0.7]),
            {"id": "metal", "type": "metal", "data": {"color": [0.9, 0.9, 0.9]}},
            {"id": "glass", "type": "dielectric", "data": {"color": [0.9, 0.9, 0.9]}},
            light("light", [25.0, 25.0, 25.0])]
    inst = [direct(rect("white", [80, 80])),
            direct(rect("light", [5, 5]), translate=[0, 20.0, 10]),
            direct(sphere("green", 4), translate=[-10, 4, 0]),
            direct(sphere("glass", 4), translate=[0, 4, 0]),
            direct(sphere("metal", 4), translate=[10, 4, 0])]
    return scene(camera([0.0, 20.0, -40.0], [0.0, 0.0, 0.0], 40.0, 40.0), [0.0, 0.0, 0.0], mats, [], inst)


def cornell(green, red, light_rgb, lights):
    mats = [lambertian("green", green), lambertian("red", red), lambertian("white", [0.73, 0.73, 0.73]),
            light("light", light_rgb)]
    prims = [WHITE_WALL, dict(id="box", **box("white", [165, 165, 165]))]
    inst = cornell_shell() + [ref("box", translate=[212.5, 82.5, 147.5], rotate=[0.0, -0.1, 0.0]), TALL_BOX]
    inst += lights + [SKIPPED_SPHERE_LIGHT]
    return scene(camera([278.0, 278.0, -750.0], [278.0, 278.0, 0.0], 40.0, 10.0), [0.0, 0.0, 0.0], mats, prims, inst)


def cornell_box():
    return cornell([0.12, 0.85, 0.05], [0.95, 0.05, 0.05], [0.6, 0.6, 0.6],
                   [direct(rect("light", [240, 230]), translate=[273, 554.0, 171])])


def cornell_box_small_lights():
    return cornell([0.12, 0.45, 0.15], [0.65, 0.05, 0.05], [15.0, 15.0, 15.0],
                   [direct(rect("light", [30, 30]), translate=[243, 554.0, 171]),
                    direct(rect("light", [30, 30]), translate=[303, 554.0, 171])])


def cornell_volume(world_rgb, iso_rgb, iso_density, box_size, fog_density, fog_transform):
    mats = [lambertian("green", [0.12, 0.45, 0.15]), lambertian("red", [0.65, 0.05, 0.05]),
            lambertian("white", [0.73, 0.73, 0.73]),
            {"id": "isotropic", "type": "isotropic", "data": {"color": iso_rgb, "density": iso_density}},
            light("light", [1.0, 1.0, 1.0])]
    prims = [WHITE_WALL, {"id": "box", "type": "box", "size": box_size},
             {"id": "fog", "type": "volume", "primitive": "box", "density": fog_density, "color": [0.9, 0.9, 0.9]}]
    skipped_box = {"skip": True}
    skipped_box.update(ref("box", translate=[212.5, 82.5, 147.5], rotate=[0.0, -0.1, 0.0]))
    inst = cornell_shell() + [skipped_box, ref("fog", **fog_transform), TALL_BOX,
                              direct(rect("light", [240, 230]), translate=[273, 554.0, 171]), SKIPPED_SPHERE_LIGHT]
    return scene(camera([278.0, 278.0, -700.0], [278.0, 278.0, 0.0], 40.0, 10.0), world_rgb, mats, prims, inst)


def cornell_box_with_volume():
    return cornell_volume([0.1, 0.1, 0.1], [0.4, 0.4, 0.4], 0.004, [165, 165, 165], 0.004,
                          dict(translate=[212.5, 82.5, 147.5], rotate=[0.0, -0.1, 0.0]))


def cornell_box_with_volume2():
    return cornell_volume([0.0, 0.0, 0.0], [0.93, 0.93, 0.93], 0.0004, [555, 555, 555], 0.001,
                          dict(translate=[277.5, 277.5, 277.5]))


def light_test():
    mats = [lambertian("white", [0.7, 0.7, 0.7])]
    for k in (1, 2, 3):
        mats.append({"id": "metal%d" % k, "type": "metal", "data": {"color": [0.9, 0.9, 0.9], "roughness": k / 10}})
    mats.append(light("light", [25.0, 25.0, 25.0]))
    prims = [dict(id="wall", **rect("white", [80, 80]))]
    inst = [ref("wall"), ref("wall", rotate=[-0.5, 0.0, 0.0], translate=[0.0, 40.0, 40.0])]
    for x, r in ((-15, 2), (-5, 1.5), (5, 1), (15, 0.5)):
        inst.append(direct(sphere("light", r), translate=[x, 20.0, 4.0]))
    inst.append(direct(rect("metal3", [40, 4.8]), rotate=[-0.0922222, 0.0, 0.0], translate=[0, 3, 0]))
    inst.append(direct(rect("metal2", [40, 4]), rotate=[-0.155, 0.0, 0.0], translate=[0, 4, 6]))
    inst.append(direct(rect("metal1", [40, 4]), rotate=[-0.2177777, 0.0, 0.0], translate=[0, 10, 8]))
    return scene(camera([0.0, 80.0, -80.0], [0.0, 0.0, 0.0], 40.0, 40.0), [0.01, 0.01, 0.01], mats, prims, inst)


def book2_final(seed=1234, n_instances=8192):
    """Synthetic "Ray Tracing: The Next Week" final scene in the reference's schema:
    exactly n_instances (a power of two => no n==1 BVH leaves, quirk Q9) instances: 64x64 ground boxes,
    one rect light, three large spheres, a glass sphere wrapping a blue medium, a global mist medium,
    a checker sphere, a Perlin sphere, and a rotated cloud of small white spheres."""
    rng = np.random.Generator(np.random.PCG64(seed))
    textures = [
        {"id": "checker_a", "type": "constant", "data": {"color": [0.2, 0.3, 0.1]}},
        {"id": "checker_b", "type": "constant", "data": {"color": [0.9, 0.9, 0.9]}},
        {"id": "checker", "type": "checker", "data": {"scale": 0.1, "odd": {"texture": "checker_a"},
                                                      "even": {"texture": "checker_b"}}},
        {"id": "noise", "type": "perlin", "data": {"scale": 0.1}},
    ]
    mats = [lambertian("ground", [0.48, 0.83, 0.53]), light("light", [7.0, 7.0, 7.0]),
            lambertian("orange", [0.7, 0.3, 0.1]),
            {"id": "glass", "type": "dielectric", "data": {"ior": 1.5}},
            {"id": "steel", "type": "metal", "data": {"color": [0.8, 0.8, 0.9], "roughness": 1.0}},
            {"id": "checker", "type": "lambertian", "data": {"texture": "checker"}},
            {"id": "noise", "type": "lambertian", "data": {"texture": "noise"}},
            lambertian("white", [0.73, 0.73, 0.73])]
    prims = [
        {"id": "unit_box", "type": "box", "material": {"id": "ground"}, "p0": [0, 0, 0], "p1": [1, 1, 1]},
        dict(id="small_sphere", **sphere("white", 10)),
        dict(id="subsurface_boundary", **sphere("glass", 70)),
        {"id": "subsurface", "type": "volume", "primitive": "subsurface_boundary", "density": 0.2,
         "color": [0.2, 0.4, 0.9]},
        dict(id="mist_boundary", **sphere("glass", 5000)),
        {"id": "mist", "type": "volume", "primitive": "mist_boundary", "density": 0.0001, "color": [1.0, 1.0, 1.0]},
    ]
    inst = []
    w = 2000.0 / 64
    for i in range(64):
        for j in range(64):
            h = float(np.round(rng.uniform(1.0, 101.0), 3))
            inst.append(ref("unit_box", scale=[w, h, w], translate=[-1000.0 + i * w, 0.0, -1000.0 + j * w]))
    inst.append(direct(rect("light", [300, 265]), translate=[273, 554.0, 279.5]))
    inst.append(direct(sphere("orange", 50), translate=[400, 400, 200]))
    inst.append(direct(sphere("glass", 50), translate=[260, 150, 45]))
    inst.append(direct(sphere("steel", 50), translate=[0, 150, 145]))
    inst.append(ref("subsurface_boundary", translate=[360, 150, 145]))
    inst.append(ref("subsurface", translate=[360, 150, 145]))
    inst.append(ref("mist"))
    inst.append(direct(sphere("checker", 100), translate=[400, 200, 400]))
    inst.append(direct(sphere("noise", 80), translate=[220, 280, 300]))
    n_small = n_instances - len(inst)
    # cloud of small spheres: uniform in a 165^3 cube, rotated 15 degrees about y, then translated.
    ang = np.deg2rad(15.0)
    c, s = np.cos(ang), np.sin(ang)
    pts = rng.uniform(0.0, 165.0, size=(n_small, 3))
    for p in pts:
        x, y, z = p
        xr, zr = c * x + s * z, -s * x + c * z
        inst.append(ref("small_sphere", translate=[float(np.round(xr - 100.0, 3)), float(np.round(y + 270.0, 3)),
                                                   float(np.round(zr + 395.0, 3))]))
    assert len(inst) == n_instances
    return scene(camera([478.0, 278.0, -600.0], [278.0, 278.0, 0.0], 40.0, 10.0), [0.0, 0.0, 0.0], mats, prims,
                 inst, textures)


SCENES = {
    "three_orbs": three_orbs,
    "cornell_box": cornell_box,
    "cornell_box_small_lights": cornell_box_small_lights,
    "cornell_box_with_volume": cornell_box_with_volume,
    "cornell_box_with_volume2": cornell_box_with_volume2,
    "light_test": light_test,
    "book2_final": book2_final,
}


def main():
    outdir = sys.argv[1] if len(sys.argv) > 1 else os.path.dirname(os.path.abspath(__file__))
    for name, fn in SCENES.items():
        with open(os.path.join(outdir, name + ".json"), "w") as f:
            json.dump(fn(), f, separators=(",", ":"))
            f.write("\n")
    print("wrote", ", ".join(SCENES))


if __name__ == "__main__":
    main()
