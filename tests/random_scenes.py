"""Seeded random scenes in the reference's JSON format: spheres, boxes and rects under random transforms (some of them
wrapped in constant media), with clusters, nested shells and coincident surfaces -- shapes of trees the shipped scenes do
not have."""
import json

import numpy as np


def random_scene(n, seed, media=True):
    rng = np.random.Generator(np.random.PCG64(seed))
    mats = [{"id": "white", "type": "lambertian", "data": {"color": [0.7, 0.7, 0.7]}},
            {"id": "metal", "type": "metal", "data": {"color": [0.8, 0.8, 0.9]}},
            {"id": "glass", "type": "dielectric", "data": {"color": [1.0, 1.0, 1.0]}},
            {"id": "fog", "type": "isotropic", "data": {"color": [0.9, 0.9, 0.9]}},
            {"id": "light", "type": "diffuse_light", "data": {"color": [15.0, 15.0, 15.0]}}]
    inst = [{"type": "direct", "primitive": {"type": "rect", "material": {"id": "light"}, "size": [6, 6]},
             "transform": {"translate": [0, 30.0, 0], "rotate": [1.0, 0, 0]}},
            {"type": "direct", "primitive": {"type": "rect", "material": {"id": "white"}, "size": [120, 120]}}]
    prims, n_media = [], 0
    centres = rng.uniform(-25, 25, size=(max(1, n // 40), 3))
    while len(inst) < n:
        kind = rng.integers(0, 10)
        c = centres[rng.integers(0, len(centres))] + rng.normal(scale=4.0, size=3)
        c[1] = abs(c[1])
        mat = ["white", "metal", "glass"][rng.integers(0, 3)]
        tr = {"translate": [float(x) for x in c]}
        if rng.random() < 0.5:
            tr["rotate"] = [float(x) for x in rng.uniform(-1, 1, size=3) * (rng.random(3) < 0.6)]
        if kind < 5:
            prim = {"type": "sphere", "material": {"id": mat}, "radius": float(rng.uniform(0.3, 2.5))}
            if rng.random() < 0.1 and len(inst) + 1 < n:  # a shell: the same sphere again, slightly smaller
                inst.append({"type": "direct", "primitive": dict(prim, radius=prim["radius"] * 0.9), "transform": tr})
        elif kind < 8:
            tr["scale"] = [float(x) for x in rng.uniform(0.5, 4.0, size=3)]
            prim = {"type": "box", "material": {"id": mat}, "size": [1.0, 1.0, 1.0]}
        else:
            prim = {"type": "rect", "material": {"id": mat}, "size": [float(x) for x in rng.uniform(1, 8, size=2)]}
            if rng.random() < 0.3:
                tr["translate"][1] = 0.0  # coincident with the floor: the tie rule decides
                tr.pop("rotate", None)
        if media and kind < 8 and n_media < 20 and rng.random() < 0.05:  # "volume" wraps a NAMED primitive
            pid = "boundary%d" % n_media
            n_media += 1
            prims.append(dict(prim, id=pid))
            prim = {"type": "volume", "primitive": pid, "density": float(rng.uniform(0.01, 0.3)),
                    "color": [0.9, 0.9, 0.9]}
        inst.append({"type": "direct", "primitive": prim, "transform": tr})
    scene = {"camera": {"look_from": [0.0, 25.0, -70.0], "look_at": [0.0, 5.0, 0.0], "fov": 40.0, "aperture": 0.0,
                        "dist_to_focus": 70.0},
             "world": {"color": [0.2, 0.3, 0.5]}, "assets": [], "textures": [], "materials": mats, "primitives": prims,
             "instances": inst[:n]}
    return json.dumps(scene)
