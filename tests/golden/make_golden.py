#!/usr/bin/env python3
"""Generates the golden vectors under tests/golden/ by running the UNMODIFIED reference (oracle/_ref/ref_driver,
compiled from /root/reference by oracle/Makefile).  The reference has no test vectors of its own,
so these are its outputs on seeded inputs:

  scene_<name>.json   ref_driver `scene`: BVH leaves (transform, primitive, bbox, material, light, n==1 duplication),
                      camera, Perlin tables                                      -> tests/test_scene_compile.py
  hits_<name>.npz     ref_driver `hit` on a seeded ray batch; instance ids re-numbered into the flat scene's
                      instance table (tests/scene_match.py)                       -> test_oracle_hit.py, test_gpu_trace.py
  color_<tag>.npz     ref_driver `color`: per-pixel mean / clamped mean / clamped M2 of integrator->color over N
                      samples, single-threaded (bit-reproducible)                 -> test_oracle_vs_reference.py,
                                                                                    test_gpu_render.py
usage: python tests/golden/make_golden.py [scene|hits|color ...]
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import raygen  # noqa: E402
import scene_match  # noqa: E402
from oracle_lib import PortOracle, RefDriver  # noqa: E402

SMALL = ["three_orbs", "cornell_box", "cornell_box_small_lights", "cornell_box_with_volume", "cornell_box_with_volume2",
         "light_test", "cornell_box_image_textures"]
ONLY = os.environ.get("GOLDEN_ONLY", "")  # regenerate only the scenes / tags containing this substring
ONLY_EXACT = os.environ.get("GOLDEN_TAG", "")  # regenerate exactly this color tag
ALL = SMALL + ["book2_final"]
W, H = 32, 18


def scene_file(name):
    return os.path.join(ROOT, "scenes", name + ".json")


def flat_of(name, w=W, h=H):
    import pathtrace_b200 as pt
    return pt.Scene.from_file(scene_file(name), w, h, device=-1)  # host-only handle: no GPU needed


def make_scene_dumps():
    for name in [n for n in SMALL if ONLY in n]:
        dump = RefDriver.scene(scene_file(name), W, H)
        with open(os.path.join(HERE, f"scene_{name}.json"), "w") as f:
            json.dump(dump, f, separators=(",", ":"))
        print("scene", name, dump["n_leaves"])


def make_hits():
    for name in [n for n in ALL if ONLY in n]:
        sc = flat_of(name)
        flat = sc.describe()
        ref_scene = RefDriver.scene(scene_file(name), W, H)
        mapping, worst = scene_match.match_leaves(ref_scene, flat)
        assert worst < 1e-6, worst
        info = sc.bvh_info()
        cam = raygen.camera_rays(ref_scene["camera"], W, H)
        rnd = raygen.random_rays(info["bounds"][:3], info["bounds"][3:], 2500, seed=101)
        first = scene_match.remap_ref_hits(RefDriver.hit(scene_file(name), cam), mapping)
        sec = raygen.secondary_rays(first, 1020, seed=102)
        rays = np.concatenate([cam, rnd, sec])
        hits = scene_match.remap_ref_hits(RefDriver.hit(scene_file(name), rays), mapping)
        np.savez_compressed(os.path.join(HERE, f"hits_{name}.npz"), rays=rays,
                            **{f: hits[f] for f in ("instance", "t", "p", "n", "u", "v")})
        print("hits", name, len(rays), "hit fraction", float((hits["instance"] >= 0).mean()))


COLOR_CASES = {
    # tag: (scene, w, h, n, clamp, config overrides)
    "cornell_box_d1": ("cornell_box", 32, 32, 2048, 3.0, dict(max_bounces=1)),
    "cornell_box_b3": ("cornell_box", 32, 32, 2048, 3.0, dict(max_bounces=3)),
    "cornell_box_rpt": ("cornell_box", 24, 24, 2048, 3.0, dict(integrator_type="recursive path tracing", max_bounces=4)),
    "three_orbs_full": ("three_orbs", 32, 18, 2048, 3.0, dict()),
    "cornell_box_small_lights_d1": ("cornell_box_small_lights", 32, 32, 2048, 3.0, dict(max_bounces=1, light_samples=1)),
    "cornell_box_with_volume_b2": ("cornell_box_with_volume", 32, 32, 2048, 3.0, dict(max_bounces=2)),
    "light_test_b2": ("light_test", 32, 18, 2048, 3.0, dict(max_bounces=2)),
    "book2_final_b2": ("book2_final", 32, 18, 2048, 3.0, dict(max_bounces=2)),
    "cornell_box_image_textures_b2": ("cornell_box_image_textures", 32, 32, 2048, 3.0, dict(max_bounces=2)),
    # the SHIPPED integrator settings (config.json:12-25: INEEPT, max_bounces 10, light_samples 4, RR on) on every
    # BASELINE scene, plus one rung each with Russian roulette off and with direct illumination only.
    # `_rep` = a second, independent replica of the reference (RNG burnt by REP_BURN draws before the scene is built):
    # the reference-vs-reference floor the block-rRMSE of the GPU is judged against.
    "cornell_box_full": ("cornell_box", 64, 64, 1024, 3.0, dict()),
    "cornell_box_full_rep": ("cornell_box", 64, 64, 1024, 3.0, dict()),
    "cornell_box_small_lights_full": ("cornell_box_small_lights", 64, 64, 1024, 3.0, dict()),
    "cornell_box_small_lights_full_rep": ("cornell_box_small_lights", 64, 64, 1024, 3.0, dict()),
    "cornell_box_with_volume_full": ("cornell_box_with_volume", 64, 64, 1024, 3.0, dict()),
    "cornell_box_with_volume_full_rep": ("cornell_box_with_volume", 64, 64, 1024, 3.0, dict()),
    "book2_final_full": ("book2_final", 64, 36, 1024, 3.0, dict()),
    "book2_final_full_rep": ("book2_final", 64, 36, 1024, 3.0, dict()),
    # the north star's literal sample count (4096 spp), for the "rRMSE @ 4096 spp next to the floor" report
    "cornell_box_4096": ("cornell_box", 64, 64, 4096, 3.0, dict()),
    "cornell_box_4096_rep": ("cornell_box", 64, 64, 4096, 3.0, dict()),
    "cornell_box_norr": ("cornell_box", 48, 48, 1024, 3.0, dict(russian_roulette=False)),
    "cornell_box_direct": ("cornell_box", 48, 48, 1024, 3.0, dict(only_direct_illumination=True)),
    "cornell_box_with_volume_norr": ("cornell_box_with_volume", 32, 32, 1024, 3.0, dict(russian_roulette=False)),
}
REP_BURN = 10_000_019  # draws discarded by the `_rep` replicas


def make_color():
    import pathtrace_b200 as pt
    for tag, (scene, w, h, n, clamp, over) in COLOR_CASES.items():
        if ONLY not in tag or (ONLY_EXACT and tag != ONLY_EXACT):
            continue
        cfgd = pt.default_config_dict(width=w, height=h, samples=n, scene=scene_file(scene), **over)
        rep = tag.endswith("_rep")
        out = RefDriver.color(cfgd, n, clamp, burn=REP_BURN if rep else 0)
        cfg_store = dict(cfgd, scene=f"scenes/{scene}.json")
        big = w * h > 32 * 32  # the larger films are stored as float32 (1e-7 relative: far below their sampling noise)
        f = (lambda a: a.astype(np.float32)) if big else (lambda a: a)
        np.savez_compressed(os.path.join(HERE, f"color_{tag}.npz"), config=json.dumps(cfg_store), n=n, clamp=clamp,
                            mean=f(out["mean"]), cmean=f(out["cmean"]), cm2=f(out["cm2"]), total_rays=out["total_rays"])
        print("color", tag, "mean radiance", float(out["mean"].mean()), "rays", out["total_rays"])


if __name__ == "__main__":
    what = sys.argv[1:] or ["scene", "hits", "color"]
    if "scene" in what:
        make_scene_dumps()
    if "hits" in what:
        make_hits()
    if "color" in what:
        make_color()
