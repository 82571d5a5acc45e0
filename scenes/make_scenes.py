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


def cornell_box_image_textures():
    """Not one of the reference's files: a Cornell box that exercises `png` textures (scene_parser.h:318-323,
    image.h) on a lambertian back wall and box (rect u, v incl. the `v` quirk of primitive.h:207) and as the world
    texture seen through the open front (integrator.h:325-335).  Paths are relative to the repository root, where
    the reference harness runs.  The light keeps a constant texture: the reference itself cannot render an
    image-textured rect light under NEE -- the shadow rays it fires from a point ON the light to the light lie in
    its plane, t = 0/0 passes rect::hit and image_texture::value indexes with int(NaN) (segfault); every textured
    rect here has x0 == z0, which keeps `v` in [0, 1] and away from the out-of-range row of image.h:26 as well.
    tests/test_gpu_render.py renders the image-LIGHT variant against the CPU restatement only."""
    textures = [{"id": "panel", "type": "png", "data": {"path": "scenes/assets/panel.png"}},
                {"id": "sky", "type": "png", "data": {"path": "scenes/assets/sky.png"}}]
    mats = [lambertian("green", [0.12, 0.85, 0.05]), lambertian("red", [0.95, 0.05, 0.05]),
            lambertian("white", [0.73, 0.73, 0.73]),
            {"id": "poster", "type": "lambertian", "data": {"texture": "panel"}},
            light("light", [0.6, 0.6, 0.6])]
    prims = [WHITE_WALL, dict(id="poster_wall", **rect("poster", [555, 555])), dict(id="box", **box("poster", [165, 165, 165]))]
    inst = [ref("white_wall", translate=[277.5, 0.0, 277.5]),
            ref("white_wall", rotate=[1.0, 0.0, 0.0], translate=[277.5, 555, 277.5]),
            ref("poster_wall", rotate=[1.5, 0, 0], translate=[277.5, 277.5, 555]),
            direct(rect("green", [555, 555], align="yz", flip=True), translate=[555, 277.5, 277.5]),
            direct(rect("red", [555, 555], align="yz"), translate=[0, 277.5, 277.5]),
            ref("box", translate=[212.5, 82.5, 147.5], rotate=[0.0, -0.1, 0.0]), TALL_BOX,
            direct(rect("light", [240, 230]), translate=[273, 554.0, 171])]
    sc = scene(camera([278.0, 278.0, -750.0], [278.0, 278.0, 0.0], 40.0, 10.0), [0.0, 0.0, 0.0], mats, prims, inst,
               textures)
    sc["world"] = {"texture": "sky"}
    return sc


def write_png(path, rgba, color_type):
    """A small PNG writer (stdlib only): rgba is H x W x 4 uint8; color_type 6 = RGBA8, 2 = RGB8.  Rows cycle through
    the five scanline filters so that a reader has to undo every one of them."""
    import struct
    import zlib
    h, w, _ = rgba.shape
    ch = 4 if color_type == 6 else 3
    rows = rgba[:, :, :ch].astype(np.int32).reshape(h, w * ch)
    raw = bytearray()
    prev = np.zeros(w * ch, dtype=np.int32)
    for y in range(h):
        cur = rows[y]
        a = np.concatenate([np.zeros(ch, np.int32), cur[:-ch]])
        c = np.concatenate([np.zeros(ch, np.int32), prev[:-ch]])
        f = y % 5
        if f == 0:
            out = cur
        elif f == 1:
            out = cur - a
        elif f == 2:
            out = cur - prev
        elif f == 3:
            out = cur - ((a + prev) >> 1)
        else:
            p = a + prev - c
            pa, pb, pc = np.abs(p - a), np.abs(p - prev), np.abs(p - c)
            pred = np.where((pa <= pb) & (pa <= pc), a, np.where(pb <= pc, prev, c))
            out = cur - pred
        raw.append(f)
        raw += bytes((out & 255).astype(np.uint8))
        prev = cur

    def chunk(tag, body):
        return struct.pack(">I", len(body)) + tag + body + struct.pack(">I", zlib.crc32(tag + body) & 0xffffffff)

    data = b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, color_type, 0, 0, 0))
    data += chunk(b"IDAT", zlib.compress(bytes(raw), 9)) + chunk(b"IEND", b"")
    with open(path, "wb") as f:
        f.write(data)


def write_assets(outdir):
    os.makedirs(os.path.join(outdir, "assets"), exist_ok=True)
    rng = np.random.Generator(np.random.PCG64(99))
    # panel: 16 x 8 colour blocks with a soft random tint; alpha falls off towards the left edge, with two holes
    h, w = 8, 16
    yy, xx = np.mgrid[0:h, 0:w]
    panel = np.zeros((h, w, 4), dtype=np.uint8)
    panel[..., 0] = 40 + 200 * ((xx // 4) % 2)
    panel[..., 1] = 60 + 180 * ((yy // 2) % 2)
    panel[..., 2] = 255 - 12 * xx
    panel[..., :3] = np.clip(panel[..., :3].astype(np.int32) + rng.integers(-20, 21, size=(h, w, 3)), 0, 255)
    panel[..., 3] = np.clip(64 + 16 * xx, 0, 255)
    panel[2, 5, 3] = 0
    panel[6, 11, 3] = 0
    write_png(os.path.join(outdir, "assets", "panel.png"), panel, 6)
    # sky: 32 x 16 RGB gradient with a bright band
    h, w = 16, 32
    yy, xx = np.mgrid[0:h, 0:w]
    sky = np.zeros((h, w, 4), dtype=np.uint8)
    sky[..., 0] = 30 + 6 * xx
    sky[..., 1] = 80 + 8 * yy
    sky[..., 2] = 230 - 5 * yy
    sky[5:7, :, :3] = 250
    sky[..., 3] = 255
    write_png(os.path.join(outdir, "assets", "sky.png"), sky, 2)


SCENES = {
    "cornell_box_image_textures": cornell_box_image_textures,
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
    write_assets(outdir)
    print("wrote", ", ".join(SCENES), "+ assets/panel.png, assets/sky.png")


if __name__ == "__main__":
    main()
